/* Build configuration for compiling the reference (read-only, /root/reference) in place.
 * Mirrors /root/reference/configure/config_core.in as CMake would emit it for
 *   -DNATIVE_SIZE=32 -DMATHBACKEND=4 -DWITH_OPENMP=OFF   (README.md:19 of the reference).
 * Test infrastructure only. */
#ifndef MKFHE_B200_ORACLE_CONFIG_CORE_H
#define MKFHE_B200_ORACLE_CONFIG_CORE_H
#define WITH_BE4
#define HAVE_INT64 1
#define MATHBACKEND 4
#define NATIVEINT 32
#endif
