"""ctypes binding of oracle/_ref/libmkfhe_ref.so -- the UNMODIFIED reference compiled in place (oracle/Makefile, `make ref`).

TEST INFRASTRUCTURE ONLY.  Imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs,
never by the product package (mkfhe_b200/).  The library is built in the CPU container from /root/reference and travels to
the GPU box as a prebuilt .so; nothing here reads /root/reference at run time.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libmkfhe_ref.so")

U32P = C.POINTER(C.c_uint32)


def available() -> bool:
    return os.path.exists(LIB_PATH)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise FileNotFoundError(f"{LIB_PATH} missing: run `make -C oracle ref` where /root/reference exists")
        L = C.CDLL(LIB_PATH)
        L.ref_create.restype = C.c_void_p
        L.ref_create.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                 C.c_double, C.c_uint32, C.c_uint32, C.c_uint32]
        L.ref_destroy.argtypes = [C.c_void_p]
        L.ref_last_error.restype = C.c_char_p
        L.ref_last_error.argtypes = [C.c_void_p]
        L.ref_get_params.argtypes = [C.c_void_p, U32P]
        L.ref_keygen.argtypes = [C.c_void_p]
        L.ref_bk_words.restype = C.c_uint64
        L.ref_bk_words.argtypes = [C.c_void_p]
        for name in ("ref_export_bk", "ref_export_pkey", "ref_export_ctnand", "ref_export_sk", "ref_export_unienc_sk", "ref_export_crs", "ref_export_unienc_sk_coef"):
            getattr(L, name).argtypes = [C.c_void_p, U32P]
        L.ref_export_mntru_F.argtypes = [C.c_void_p, U32P, C.c_int]
        L.ref_export_ksk_mntru.argtypes = [C.c_void_p, U32P, C.POINTER(C.c_uint64)]
        L.ref_export_ksk_mklwe.argtypes = [C.c_void_p, U32P, U32P]
        L.ref_encrypt.argtypes = [C.c_void_p, C.c_int, U32P, U32P]
        L.ref_decrypt.argtypes = [C.c_void_p, U32P, C.c_uint32, C.POINTER(C.c_int)]
        L.ref_eval_acc.argtypes = [C.c_void_p, U32P, U32P, C.c_uint32]
        L.ref_eval_gate.argtypes = [C.c_void_p, U32P, C.c_uint32, U32P, C.c_uint32, U32P, U32P]
        L.ref_ntt.argtypes = [C.c_void_p, U32P, C.c_uint64, C.c_int]
        L.ref_decompose.argtypes = [C.c_void_p, U32P, U32P]
        L.ref_bench_gates.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_double)]
        L.ref_time_gate.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_double)]
        L.ref_use_accumulator_b200.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.ref_batch_b200_load.argtypes = [C.c_void_p, C.c_char_p]
        L.ref_batch_b200_save_keyfile.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        L.ref_batch_b200_load_keyfile.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        L.ref_batch_b200_ct_roundtrip.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_uint8), C.c_uint64, U32P, U32P, U32P, U32P,
                                                  C.POINTER(C.c_int)]
        L.ref_batch_b200_gates.argtypes = [C.c_void_p, U32P, U32P, U32P, U32P, U32P, U32P, C.c_uint64, C.POINTER(C.c_int)]
        L.ref_batch_b200_errors.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        L.ref_batch_b200_circuit.argtypes = [C.c_void_p, U32P, U32P, C.POINTER(C.c_uint8), C.c_uint64, U32P, U32P, U32P, C.c_uint64,
                                             U32P, C.POINTER(C.c_int)]
        _lib = L
    return _lib


def _p(a: np.ndarray):
    assert a.dtype == np.uint32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(U32P)


@dataclass
class RefParams:
    k: int
    n: int
    N: int
    q: int
    qKS: int
    Q: int
    baseG: int
    digitsG: int
    baseKS: int
    dKS: int
    psi: int
    scheme: int  # 0 = MKNTRU (XZW), 1 = MKNTRU_LWE (XZW_B)

    @property
    def ell(self) -> int:
        return self.digitsG - 1


# Rows of the reference's parameter table that the configs use (src/binfhe/lib/binfhecontext.cpp:137-144):
#   name: (scheme, k, numberBits, cyclOrder, n, q, qKS, stdDev, baseKS, baseG, baseR)
PARAMSETS = {
    "STD100_MKNTRU": (0, 2, 27, 4096, 560, 45181, 45181, 0.75, 32, 1 << 9, 32),
    "STD100_MKNTRU_2": (0, 4, 27, 4096, 560, 45181, 45181, 0.75, 32, 1 << 9, 32),
    "STD100_MKNTRU_3": (0, 8, 27, 4096, 560, 45181, 45181, 0.75, 32, 1 << 9, 32),
    "STD100_MKNTRU_4": (0, 16, 27, 4096, 560, 45181, 45181, 0.75, 32, 1 << 9, 32),
    "STD100_MKNTRU_LWE": (1, 2, 27, 4096, 500, 32749, 32749, 1.9, 32, 1 << 9, 2),
    "STD100_MKNTRU_LWE_2": (1, 4, 27, 4096, 500, 32749, 32749, 1.9, 32, 1 << 9, 2),
    "STD128_MKNTRU": (0, 2, 27, 4096, 765, 45181, 45181, 0.5, 32, 1 << 7, 32),
}


class Ref:
    """One reference context: parameters + (after keygen) secret keys, bootstrapping keys, ctNAND."""

    def __init__(self, scheme, k, numberBits, cyclOrder, n, q, qKS, stdDev, baseKS, baseG, baseR):
        self.L = lib()
        self.h = self.L.ref_create(scheme, k, numberBits, cyclOrder, n, q, qKS, stdDev, baseKS, baseG, baseR)
        if not self.h:
            raise RuntimeError("ref_create failed")
        out = np.zeros(11, dtype=np.uint32)
        self.L.ref_get_params(self.h, _p(out))
        v = [int(x) for x in out]
        self.p = RefParams(k=v[0], n=v[1], N=v[2], q=v[3], qKS=v[4], Q=v[5], baseG=v[6], digitsG=v[7], baseKS=v[8],
                           dKS=v[9], psi=v[10], scheme=scheme)

    @classmethod
    def from_paramset(cls, name: str) -> "Ref":
        return cls(*PARAMSETS[name])

    def close(self):
        if self.h:
            self.L.ref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, what):
        if rc != 0:
            raise RuntimeError(f"{what}: {self.L.ref_last_error(self.h).decode()}")

    def keygen(self):
        self._ck(self.L.ref_keygen(self.h), "ref_keygen")

    # ---- exports ----
    def bk(self) -> np.ndarray:
        a = np.zeros(self.L.ref_bk_words(self.h), dtype=np.uint32)
        self._ck(self.L.ref_export_bk(self.h, _p(a)), "ref_export_bk")
        return a

    def pkey(self) -> np.ndarray:
        a = np.zeros(self.p.k * self.p.ell * self.p.N, dtype=np.uint32)
        self._ck(self.L.ref_export_pkey(self.h, _p(a)), "ref_export_pkey")
        return a

    def ksk_mntru(self):
        a = np.zeros(self.p.k * self.p.N * self.p.dKS * self.p.n, dtype=np.uint32)
        bad = C.c_uint64(0)
        self._ck(self.L.ref_export_ksk_mntru(self.h, _p(a), C.byref(bad)), "ref_export_ksk_mntru")
        return a, int(bad.value)

    def ksk_mklwe(self):
        rows = self.p.k * self.p.N * self.p.baseKS * self.p.dKS
        A = np.zeros(rows * self.p.n, dtype=np.uint32)
        B = np.zeros(rows, dtype=np.uint32)
        self._ck(self.L.ref_export_ksk_mklwe(self.h, _p(A), _p(B)), "ref_export_ksk_mklwe")
        return A, B

    def ctnand(self) -> np.ndarray:
        a = np.zeros(self.p.k * self.p.n, dtype=np.uint32)
        self._ck(self.L.ref_export_ctnand(self.h, _p(a)), "ref_export_ctnand")
        return a

    def sk(self) -> np.ndarray:
        a = np.zeros(self.p.k * self.p.n, dtype=np.uint32)
        self._ck(self.L.ref_export_sk(self.h, _p(a)), "ref_export_sk")
        return a

    def unienc_sk(self) -> np.ndarray:
        """f_u in EVALUATION form, [k][N] (UniEncBTKey::f)."""
        a = np.zeros(self.p.k * self.p.N, dtype=np.uint32)
        self._ck(self.L.ref_export_unienc_sk(self.h, _p(a)), "ref_export_unienc_sk")
        return a.reshape(self.p.k, self.p.N)

    def unienc_sk_coef(self) -> np.ndarray:
        """f_u in COEFFICIENT form mod Q, [k][N] (UniEncBTKey::fvec)."""
        a = np.zeros(self.p.k * self.p.N, dtype=np.uint32)
        self._ck(self.L.ref_export_unienc_sk_coef(self.h, _p(a)), "ref_export_unienc_sk_coef")
        return a.reshape(self.p.k, self.p.N)

    def mntru_F(self, inverse: bool) -> np.ndarray:
        """MNTRU secret matrices F (or F^-1) mod qKS, [k][n][n]."""
        a = np.zeros(self.p.k * self.p.n * self.p.n, dtype=np.uint32)
        self._ck(self.L.ref_export_mntru_F(self.h, _p(a), int(inverse)), "ref_export_mntru_F")
        return a.reshape(self.p.k, self.p.n, self.p.n)

    def crs(self) -> np.ndarray:
        """CRS in EVALUATION form, [ell][N]."""
        a = np.zeros(self.p.ell * self.p.N, dtype=np.uint32)
        self._ck(self.L.ref_export_crs(self.h, _p(a)), "ref_export_crs")
        return a.reshape(self.p.ell, self.p.N)

    # ---- scheme ops ----
    def encrypt(self, m: int):
        ct = np.zeros(self.p.k * self.p.n, dtype=np.uint32)
        b = C.c_uint32(0)
        self._ck(self.L.ref_encrypt(self.h, int(m), _p(ct), C.byref(b)), "ref_encrypt")
        return ct, int(b.value)

    def decrypt(self, ct: np.ndarray, b: int = 0) -> int:
        m = C.c_int(0)
        ct = np.ascontiguousarray(ct, dtype=np.uint32)
        self._ck(self.L.ref_decrypt(self.h, _p(ct), int(b), C.byref(m)), "ref_decrypt")
        return int(m.value)

    def eval_acc(self, acc: np.ndarray, rot: np.ndarray, rot_mod: int) -> np.ndarray:
        a = np.ascontiguousarray(acc, dtype=np.uint32).copy()
        r = np.ascontiguousarray(rot, dtype=np.uint32)
        self._ck(self.L.ref_eval_acc(self.h, _p(a), _p(r), int(rot_mod)), "ref_eval_acc")
        return a

    def eval_gate(self, ct1, b1, ct2, b2):
        out = np.zeros(self.p.k * self.p.n, dtype=np.uint32)
        bout = C.c_uint32(0)
        ct1 = np.ascontiguousarray(ct1, dtype=np.uint32)
        ct2 = np.ascontiguousarray(ct2, dtype=np.uint32)
        self._ck(self.L.ref_eval_gate(self.h, _p(ct1), int(b1), _p(ct2), int(b2), _p(out), C.byref(bout)), "ref_eval_gate")
        return out, int(bout.value)

    def ntt(self, polys: np.ndarray, inverse: bool) -> np.ndarray:
        a = np.ascontiguousarray(polys, dtype=np.uint32).copy()
        batch = a.size // self.p.N
        self._ck(self.L.ref_ntt(self.h, _p(a), batch, 1 if inverse else 0), "ref_ntt")
        return a

    def decompose(self, poly: np.ndarray) -> np.ndarray:
        a = np.ascontiguousarray(poly, dtype=np.uint32)
        out = np.zeros(self.p.ell * self.p.N, dtype=np.uint32)
        self._ck(self.L.ref_decompose(self.h, _p(a), _p(out)), "ref_decompose")
        return out.reshape(self.p.ell, self.p.N)

    def bench_gates(self, gates: int, threads: int) -> float:
        s = C.c_double(0.0)
        self._ck(self.L.ref_bench_gates(self.h, gates, threads, C.byref(s)), "ref_bench_gates")
        return float(s.value)

    def time_gate(self, reps: int = 5):
        """Wall milliseconds of `reps` single EvalBinGate(NAND) calls on one thread (the reference example's own measurement)."""
        ms = (C.c_double * reps)()
        self._ck(self.L.ref_time_gate(self.h, reps, ms), "ref_time_gate")
        return [float(x) for x in ms]

    def use_accumulator_b200(self, libpath: str, restore: bool = False):
        self._ck(self.L.ref_use_accumulator_b200(self.h, libpath.encode(), 1 if restore else 0), "ref_use_accumulator_b200")

    # ---- integration/binfhe_batch_b200.h: EvalBinGateBatch over the reference's own key and ciphertext objects ----
    def batch_b200_load(self, libpath: str):
        self._ck(self.L.ref_batch_b200_load(self.h, libpath.encode()), "ref_batch_b200_load")

    def batch_b200_save_keyfile(self, libpath: str, path: str):
        """BinFHEBatchB200::SaveKeyFile over this context's UniEncBTKey (+ ctNAND): host only, no GPU."""
        self._ck(self.L.ref_batch_b200_save_keyfile(self.h, libpath.encode(), os.fsencode(path)), "ref_batch_b200_save_keyfile")

    def batch_b200_load_keyfile(self, libpath: str, path: str):
        """A fresh BinFHEBatchB200 created from the key file alone (replaces the evaluator batch_b200_load made)."""
        self._ck(self.L.ref_batch_b200_load_keyfile(self.h, libpath.encode(), os.fsencode(path)), "ref_batch_b200_load_keyfile")

    def batch_b200_ct_roundtrip(self, path: str, bits):
        """Encrypt `bits` with the reference, SaveCiphertexts, LoadCiphertexts*: (table in, b in, table out, b out, Decrypt of out)."""
        bits = np.ascontiguousarray(bits, dtype=np.uint8)
        n = len(bits)
        fin, fout = (np.zeros((n, self.p.k, self.p.n), dtype=np.uint32) for _ in range(2))
        bin_, bout = (np.zeros(n, dtype=np.uint32) for _ in range(2))
        dec = (C.c_int * n)()
        self._ck(self.L.ref_batch_b200_ct_roundtrip(self.h, os.fsencode(path), bits.ctypes.data_as(C.POINTER(C.c_uint8)), n, _p(fin),
                                                    _p(bin_), _p(fout), _p(bout), dec), "ref_batch_b200_ct_roundtrip")
        return fin, bin_, fout, bout, np.array(list(dec), dtype=np.int64)

    def batch_b200_gates(self, ct1, b1, ct2, b2):
        """ct1, ct2: [B][k][n]; b1, b2: [B] (MK-LWE; ignored for matrix-NTRU).  Returns (out [B][k][n], bout [B], the reference's
        Decrypt of every output)."""
        ct1 = np.ascontiguousarray(ct1, dtype=np.uint32)
        ct2 = np.ascontiguousarray(ct2, dtype=np.uint32)
        B = ct1.shape[0]
        b1 = np.ascontiguousarray(b1 if b1 is not None else np.zeros(B), dtype=np.uint32)
        b2 = np.ascontiguousarray(b2 if b2 is not None else np.zeros(B), dtype=np.uint32)
        out = np.zeros_like(ct1)
        bout = np.zeros(B, dtype=np.uint32)
        dec = (C.c_int * B)()
        self._ck(self.L.ref_batch_b200_gates(self.h, _p(ct1), _p(b1), _p(ct2), _p(b2), _p(out), _p(bout), B, dec), "ref_batch_b200_gates")
        return out, bout, np.array(list(dec), dtype=np.int64)

    def batch_b200_errors(self) -> int:
        m = C.c_int(0)
        self._ck(self.L.ref_batch_b200_errors(self.h, C.byref(m)), "ref_batch_b200_errors")
        return int(m.value)

    def batch_b200_circuit(self, wires, wb, n_inputs, gates):
        """BinFHEBatchB200::EvalCircuit: wires [W][k][n] with the first n_inputs rows set, gates [G][3].  Returns (wires, b,
        levels, the reference's Decrypt of every wire)."""
        w = np.ascontiguousarray(wires, dtype=np.uint32).copy()
        W = w.shape[0]
        b = np.ascontiguousarray(wb if wb is not None else np.zeros(W), dtype=np.uint32).copy()
        g = np.ascontiguousarray(gates, dtype=np.uint32).reshape(-1, 3)
        i1, i2, o = (np.ascontiguousarray(g[:, j]) for j in range(3))
        isin = (C.c_uint8 * W)(*([1] * n_inputs + [0] * (W - n_inputs)))
        lv = np.zeros(1, dtype=np.uint32)
        dec = (C.c_int * W)()
        self._ck(self.L.ref_batch_b200_circuit(self.h, _p(w), _p(b), isin, W, _p(i1), _p(i2), _p(o), len(g), _p(lv), dec),
                 "ref_batch_b200_circuit")
        return w, b, int(lv[0]), np.array(list(dec), dtype=np.int64)
