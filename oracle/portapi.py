"""ctypes binding of oracle/_build/libmkfhe_oracle.so -- the plain-C restatement (oracle/mkfhe_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
`build()` compiles the library with gcc when it is missing or stale (seconds; works on the GPU box too).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "mkfhe_oracle.c")
HDR = os.path.join(HERE, "mkfhe_oracle.h")
LIB_PATH = os.path.join(HERE, "_build", "libmkfhe_oracle.so")

U32P = C.POINTER(C.c_uint32)


class OrcParams(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in
                ("k", "N", "Q", "psi", "q", "qKS", "n", "baseG_log2", "digitsG", "baseKS", "dKS", "variant", "scheme")]


def build(force: bool = False) -> str:
    stale = (not os.path.exists(LIB_PATH)) or any(os.path.getmtime(f) > os.path.getmtime(LIB_PATH) for f in (SRC, HDR))
    if force or stale:
        os.makedirs(os.path.dirname(LIB_PATH), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-std=c11", "-o", LIB_PATH, SRC, "-lm", "-pthread"])
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        L.orc_first_prime.restype = C.c_uint32
        L.orc_first_prime.argtypes = [C.c_uint32, C.c_uint32]
        L.orc_previous_prime.restype = C.c_uint32
        L.orc_previous_prime.argtypes = [C.c_uint32, C.c_uint32]
        L.orc_root_of_unity.restype = C.c_uint32
        L.orc_root_of_unity.argtypes = [C.c_uint32, C.c_uint32]
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.POINTER(OrcParams)]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_ntt.argtypes = [C.c_void_p, U32P, C.c_int]
        L.orc_decompose.argtypes = [C.c_void_p, U32P, U32P]
        L.orc_monomial.argtypes = [C.c_void_p, C.c_uint32, U32P]
        L.orc_eval_acc.argtypes = [C.c_void_p, U32P, U32P, U32P, U32P, C.c_uint32]
        L.orc_eval_nand.argtypes = [C.c_void_p, U32P, U32P, U32P, U32P, U32P, U32P, C.c_uint32, U32P, C.c_uint32, U32P,
                                    U32P, U32P]
        L.orc_round_qQ.restype = C.c_uint32
        L.orc_round_qQ.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32]
        L.orc_transpose_intt.argtypes = [C.c_void_p, U32P]
        L.orc_keyswitch_mntru.argtypes = [C.c_void_p, U32P, U32P, U32P]
        L.orc_keyswitch_mklwe.argtypes = [C.c_void_p, U32P, U32P, U32P, C.c_uint32, U32P, U32P]
        L.orc_decrypt.restype = C.c_int
        L.orc_decrypt.argtypes = [C.c_void_p, U32P, U32P, C.c_uint32]
        L.orc_keygen_evk.argtypes = [C.c_void_p, U32P, U32P, U32P, U32P, C.c_int, C.c_int, U32P]
        L.orc_keygen_pkey.argtypes = [C.c_void_p, U32P, U32P, U32P, U32P]
        L.orc_keygen_ksk_mntru.argtypes = [C.c_void_p, U32P, U32P, U32P, U32P]
        L.orc_keygen_ksk_mklwe.argtypes = [C.c_void_p, U32P, U32P, U32P, U32P, U32P]
        L.orc_bench_gates.restype = C.c_double
        L.orc_bench_gates.argtypes = [C.c_void_p, U32P, U32P, U32P, U32P, U32P, U32P, C.c_uint32, U32P, C.c_uint32,
                                      C.c_uint32, C.c_uint32]
        _lib = L
    return _lib


def _p(a):
    if a is None:
        return None
    assert a.dtype == np.uint32 and a.flags["C_CONTIGUOUS"], (a.dtype, a.flags)
    return a.ctypes.data_as(U32P)


def _c(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.uint32)


def derive_modulus(number_bits: int, cycl_order: int):
    """Q and psi as the reference derives them (binfhecontext.cpp:157-158, ilparams root of unity)."""
    L = lib()
    Q = L.orc_previous_prime(L.orc_first_prime(number_bits, cycl_order), cycl_order)
    return int(Q), int(L.orc_root_of_unity(cycl_order, Q))


class Oracle:
    def __init__(self, *, k, N, Q, psi, q, qKS, n, baseG_log2, digitsG, baseKS, dKS, variant, scheme):
        self.L = lib()
        self.params = OrcParams(k, N, Q, psi, q, qKS, n, baseG_log2, digitsG, baseKS, dKS, variant, scheme)
        self.h = self.L.orc_create(C.byref(self.params))
        self.k, self.N, self.n, self.ell = k, N, n, digitsG - 1

    def __del__(self):
        try:
            if self.h:
                self.L.orc_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def ntt(self, polys, inverse: bool):
        a = _c(polys).copy().reshape(-1, self.N)
        for row in a:
            self.L.orc_ntt(self.h, _p(row), 1 if inverse else 0)
        return a.reshape(np.shape(polys))

    def decompose(self, poly):
        out = np.zeros((self.ell, self.N), dtype=np.uint32)
        self.L.orc_decompose(self.h, _p(_c(poly)), _p(out))
        return out

    def monomial(self, m: int):
        out = np.zeros(self.N, dtype=np.uint32)
        self.L.orc_monomial(self.h, m, _p(out))
        return out

    def eval_acc(self, bk, pkey, acc, rot, rot_mod: int):
        a = _c(acc).copy()
        self.L.orc_eval_acc(self.h, _p(_c(bk)), _p(_c(pkey)), _p(a), _p(_c(rot)), int(rot_mod))
        return a

    def eval_nand(self, bk, pkey, kskA, kskB, ctnand, ct1, b1, ct2, b2, trace=False):
        out = np.zeros(self.k * self.n, dtype=np.uint32)
        bout = C.c_uint32(0)
        tr = np.zeros(self.k * self.N, dtype=np.uint32) if trace else None
        self.L.orc_eval_nand(self.h, _p(_c(bk)), _p(_c(pkey)), _p(_c(kskA)), _p(_c(kskB)), _p(_c(ctnand)), _p(_c(ct1)),
                             int(b1), _p(_c(ct2)), int(b2), _p(out), C.byref(bout), _p(tr))
        return (out, int(bout.value), tr) if trace else (out, int(bout.value))

    def round_qQ(self, v, q, Q):
        return int(self.L.orc_round_qQ(int(v), int(q), int(Q)))

    def transpose_intt(self, poly):
        a = _c(poly).copy()
        self.L.orc_transpose_intt(self.h, _p(a))
        return a

    def keyswitch_mntru(self, ksk, ctQN):
        out = np.zeros(self.k * self.n, dtype=np.uint32)
        self.L.orc_keyswitch_mntru(self.h, _p(_c(ksk)), _p(_c(ctQN)), _p(out))
        return out

    def keyswitch_mklwe(self, A, B, ctQN, b):
        out = np.zeros(self.k * self.n, dtype=np.uint32)
        bout = C.c_uint32(0)
        self.L.orc_keyswitch_mklwe(self.h, _p(_c(A)), _p(_c(B)), _p(_c(ctQN)), int(b), _p(out), C.byref(bout))
        return out, int(bout.value)

    def decrypt(self, sk, ct, b=0) -> int:
        return int(self.L.orc_decrypt(self.h, _p(_c(sk)), _p(_c(ct)), int(b)))

    def keygen_evk(self, sinv, crs, r, e, m: int, kdm: bool):
        """One evk [ell][2][N] from the randomness r [N], e [ell][2][N] (coefficient residues): KeyGenXZW / KDMKeyGenXZW."""
        out = np.zeros((self.ell, 2, self.N), dtype=np.uint32)
        self.L.orc_keygen_evk(self.h, _p(_c(sinv)), _p(_c(crs)), _p(_c(r)), _p(_c(e)), int(m), int(bool(kdm)), _p(out))
        return out

    def keygen_pkey(self, s, crs, e):
        out = np.zeros((self.ell, self.N), dtype=np.uint32)
        self.L.orc_keygen_pkey(self.h, _p(_c(s)), _p(_c(crs)), _p(_c(e)), _p(out))
        return out

    def keygen_ksk_mntru(self, s_q, e, finv):
        """one party: s_q [N], e [N*dKS][n], finv [n][n] (all mod qKS) -> KSK digit-1 slice [N*dKS][n]"""
        out = np.zeros((self.N * self.params.dKS, self.n), dtype=np.uint32)
        self.L.orc_keygen_ksk_mntru(self.h, _p(_c(s_q)), _p(_c(e)), _p(_c(finv)), _p(out))
        return out

    def keygen_ksk_mklwe(self, A, e, s, sN_q):
        """one party: A [N*baseKS*dKS][n], e [rows], s [n], sN_q [N] -> B [rows]"""
        out = np.zeros(self.N * self.params.baseKS * self.params.dKS, dtype=np.uint32)
        self.L.orc_keygen_ksk_mklwe(self.h, _p(_c(A)), _p(_c(e)), _p(_c(s)), _p(_c(sN_q)), _p(out))
        return out

    def bench_gates(self, bk, pkey, kskA, kskB, ctnand, ct1, b1, ct2, b2, gates, threads) -> float:
        return float(self.L.orc_bench_gates(self.h, _p(_c(bk)), _p(_c(pkey)), _p(_c(kskA)), _p(_c(kskB)), _p(_c(ctnand)),
                                            _p(_c(ct1)), int(b1), _p(_c(ct2)), int(b2), int(gates), int(threads)))
