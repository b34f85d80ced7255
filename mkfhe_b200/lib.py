"""ctypes binding of libmkfhe_b200.so (the C ABI declared in include/mkfhe_b200.h).

The product path: there is NO CPU fallback.  Importing this module never computes anything; creating a context without a
CUDA device raises MkfheError (MKFHE_ENODEV), and a missing shared library raises ImportError with the build command.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MKFHE_B200_LIB") or os.path.join(HERE, "libmkfhe_b200.so")  # (override: experimental builds, tools/gpu_r2.sh)

U32P = C.POINTER(C.c_uint32)

MKFHE_VARIANT_XZW, MKFHE_VARIANT_XZW_B = 0, 1
MKFHE_SCHEME_MNTRU, MKFHE_SCHEME_MKLWE = 0, 1
MKFHE_IMAGE_BK, MKFHE_IMAGE_PKEY, MKFHE_IMAGE_KSK, MKFHE_IMAGE_CTNAND = 0, 1, 2, 3
CUDA_STREAM_LEGACY, CUDA_STREAM_PER_THREAD = 1, 2  # the driver's own handles (cudaStreamLegacy, cudaStreamPerThread)
ERRORS = {-1: "MKFHE_EINVAL", -2: "MKFHE_ENODEV", -3: "MKFHE_ECUDA", -4: "MKFHE_ENOKEY", -5: "MKFHE_ENOMEM"}

# every symbol include/mkfhe_b200.h declares (tests/test_cabi.py checks the header against this list and the .so)
SYMBOLS = [
    "mkfhe_ctx_create", "mkfhe_ctx_destroy", "mkfhe_last_error", "mkfhe_ctx_set_stream", "mkfhe_ctx_synchronize",
    "mkfhe_load_bk", "mkfhe_load_pkey", "mkfhe_load_ksk_mntru", "mkfhe_load_ksk_mklwe", "mkfhe_load_ctnand",
    "mkfhe_key_image", "mkfhe_key_image_commit",
    "mkfhe_eval_acc", "mkfhe_eval_acc_dev", "mkfhe_eval_nand", "mkfhe_eval_nand_dev", "mkfhe_ntt", "mkfhe_ntt_dev",
    "mkfhe_load_sk", "mkfhe_decrypt", "mkfhe_decrypt_dev", "mkfhe_eval_circuit", "mkfhe_eval_circuit_dev", "mkfhe_circuit_levels", "mkfhe_eval_circuit_dev_static", "mkfhe_circuit_static_words", "mkfhe_static_scratch_words", "mkfhe_keygen_bk", "mkfhe_keygen_pkey", "mkfhe_keygen_ksk_mntru", "mkfhe_keygen_ksk_mklwe", "mkfhe_encrypt_mntru", "mkfhe_encrypt_mklwe", "mkfhe_keyfile_write", "mkfhe_keyfile_info", "mkfhe_keyfile_load", "mkfhe_ctfile_write", "mkfhe_ctfile_info", "mkfhe_ctfile_read",
    "mkfhe_decompose", "mkfhe_last_gate_timing", "mkfhe_last_gate_sms", "mkfhe_launch_count", "mkfhe_device_bytes", "mkfhe_version", "mkfhe_measure_int32_peak", "mkfhe_measure_butterfly_peak",
]


class MkfheError(RuntimeError):
    def __init__(self, code: int, what: str, text: str):
        super().__init__(f"{what}: {ERRORS.get(code, code)}: {text}")
        self.code = code


class Params(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in ("k", "N", "Q", "psi", "q", "qKS", "n", "baseG_log2", "digitsG", "baseKS", "dKS",
                                           "variant", "scheme")] + [("device", C.c_int32)]


_lib = None


def load():
    """Load libmkfhe_b200.so; raises ImportError when it has not been built (python -m mkfhe_b200.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} not found: build the CUDA library with `python -m mkfhe_b200.build` "
                          "(nvcc, sm_100a); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
    L.mkfhe_ctx_create.argtypes = [C.POINTER(Params), C.POINTER(vp)]
    L.mkfhe_ctx_destroy.argtypes = [vp]
    L.mkfhe_ctx_destroy.restype = None
    L.mkfhe_last_error.argtypes = [vp]
    L.mkfhe_last_error.restype = C.c_char_p
    L.mkfhe_ctx_set_stream.argtypes = [vp, vp]
    L.mkfhe_ctx_synchronize.argtypes = [vp]
    for name in ("mkfhe_load_bk", "mkfhe_load_pkey", "mkfhe_load_ksk_mntru", "mkfhe_load_ctnand", "mkfhe_load_sk"):
        getattr(L, name).argtypes = [vp, vp, u64]
    L.mkfhe_load_ksk_mklwe.argtypes = [vp, vp, u64, vp, u64]
    L.mkfhe_key_image.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(u64)]
    L.mkfhe_key_image_commit.argtypes = [vp, i32]
    L.mkfhe_eval_acc.argtypes = [vp, vp, vp, u32, vp, u64]
    L.mkfhe_eval_acc_dev.argtypes = [vp, vp, vp, u32, vp, u64]
    L.mkfhe_eval_nand.argtypes = [vp, vp, vp, vp, vp, vp, vp, u64]
    L.mkfhe_eval_nand_dev.argtypes = [vp, vp, vp, vp, vp, vp, vp, u64]
    L.mkfhe_eval_circuit.argtypes = [vp, vp, vp, u64, vp, vp, vp, u64, vp]
    L.mkfhe_eval_circuit_dev.argtypes = [vp, vp, vp, u64, vp, vp, vp, u64, vp]
    L.mkfhe_circuit_levels.argtypes = [u64, vp, vp, vp, u64, vp]
    L.mkfhe_circuit_levels.restype = C.c_int64
    L.mkfhe_eval_circuit_dev_static.argtypes = [vp, vp, vp, u64, vp, vp, vp, u64, vp]
    L.mkfhe_circuit_static_words.argtypes = [C.POINTER(Params), u64, u64]
    L.mkfhe_circuit_static_words.restype = u64
    L.mkfhe_static_scratch_words.argtypes = []
    L.mkfhe_static_scratch_words.restype = u64
    L.mkfhe_decrypt.argtypes = [vp, vp, vp, u32, vp, u64]
    L.mkfhe_decrypt_dev.argtypes = [vp, vp, vp, u32, vp, u64]
    L.mkfhe_keygen_bk.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32]
    L.mkfhe_keygen_pkey.argtypes = [vp, vp, vp, vp, vp, i32]
    L.mkfhe_keygen_ksk_mntru.argtypes = [vp, vp, vp, vp, vp, i32]
    L.mkfhe_keygen_ksk_mklwe.argtypes = [vp, vp, vp, vp, vp, vp, i32]
    L.mkfhe_encrypt_mntru.argtypes = [vp, vp, vp, vp, u32, vp, u64]
    L.mkfhe_encrypt_mklwe.argtypes = [vp, vp, vp, vp, vp, u32, vp, u64]
    L.mkfhe_keyfile_write.argtypes = [C.c_char_p, C.POINTER(Params), vp, u64, vp, u64, vp, u64, vp, u64, vp, u64]
    L.mkfhe_keyfile_info.argtypes = [C.c_char_p, C.POINTER(Params), C.POINTER(C.c_uint64)]
    L.mkfhe_keyfile_load.argtypes = [vp, C.c_char_p]
    L.mkfhe_ctfile_write.argtypes = [C.c_char_p, C.POINTER(Params), vp, vp, u64]
    L.mkfhe_ctfile_info.argtypes = [C.c_char_p, C.POINTER(Params), C.POINTER(u64)]
    L.mkfhe_ctfile_read.argtypes = [C.c_char_p, C.POINTER(Params), vp, vp, u64, C.POINTER(u64)]
    L.mkfhe_ntt.argtypes = [vp, vp, u64, i32]
    L.mkfhe_ntt_dev.argtypes = [vp, vp, u64, i32]
    L.mkfhe_decompose.argtypes = [vp, vp, vp, u64]
    L.mkfhe_last_gate_timing.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    L.mkfhe_last_gate_sms.argtypes = [vp]
    L.mkfhe_last_gate_sms.restype = C.c_uint32
    L.mkfhe_launch_count.argtypes = [vp]
    L.mkfhe_launch_count.restype = u64
    L.mkfhe_device_bytes.argtypes = [vp]
    L.mkfhe_device_bytes.restype = u64
    L.mkfhe_version.restype = C.c_char_p
    L.mkfhe_measure_int32_peak.argtypes = [vp, C.POINTER(C.c_double)]
    L.mkfhe_measure_butterfly_peak.argtypes = [vp, C.POINTER(C.c_double)]
    _lib = L
    return L
