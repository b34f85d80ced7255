"""mkfhe_b200 -- B200-native multi-key accumulator bootstrapping (the hot path of SKLC-FHE/MKFHE) behind a C ABI.

    from mkfhe_b200 import BinFHEContext, MKNTRU, MKNTRU_LWE, NAND

The package holds the CUDA library (csrc/, built in-tree to libmkfhe_b200.so by `python -m mkfhe_b200.build`), its ctypes
binding (lib.py) and the host-side mirror of the reference's context API for this path (context.py).  No CPU fallback.
"""
from .context import CircuitLevels, KeyFileInfo, LoadCiphertexts, SaveCiphertexts, SaveKeyFile  # noqa: F401
from .context import (MKNTRU, MKNTRU_B, MKNTRU_LWE, NAND, PARAMSETS, BinFHEContext, BinFHEContextParams, config_error,  # noqa: F401
                      derive_params)
from .lib import MkfheError, Params  # noqa: F401

__version__ = "0.1.0"
