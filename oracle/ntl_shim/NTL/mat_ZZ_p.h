// see mkfhe_ntl_shim.h (test infrastructure; keygen-only stand-in for NTL)
#include "mkfhe_ntl_shim.h"
