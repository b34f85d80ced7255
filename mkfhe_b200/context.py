"""Host-side mirror of the reference's user API for the multi-key bootstrapping path.

Names, argument meaning and error behaviour follow `BinFHEContext` (src/binfhe/include/binfhecontext.h:94-338 of the
reference) for the calls on the path: GenerateBinFHEContext(paramset, method) -> parameters; MKBTKeyGen produces keys on the
host in the reference -- here the caller hands the reference's key bytes to `LoadKeys`; EvalBinGate(NAND, ct1, ct2) runs the
GPU path.  Everything that computes goes through the C ABI (mkfhe_b200/lib.py); this file holds no arithmetic except the
derivation of Q and psi (binfhecontext.cpp:157-158), which is host-side parameter logic.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from math import ceil, log

import numpy as np

from . import lib as _lib
from .lib import MkfheError, Params

# BINFHE_METHOD values on this path (src/binfhe/include/binfhe-constants.h)
MKNTRU, MKNTRU_B, MKNTRU_LWE = "MKNTRU", "MKNTRU_B", "MKNTRU_LWE"
NAND = "NAND"


class config_error(ValueError):
    """mirror of lbcrypto::config_error (src/core/include/utils/exception.h:101-162)"""


@dataclass(frozen=True)
class BinFHEContextParams:
    """One row of the reference's parameter table (binfhecontext.cpp:91-146): the MK rows."""
    numUser: int
    numberBits: int
    cyclOrder: int
    latticeParam: int
    mod: int
    modKS: int
    stdDev: float
    baseKS: int
    gadgetBase: int
    baseRK: int


STD_DEV, STD_NTRU, STD_NTRU2 = 1.9, 0.5, 0.75
PARAMSETS = {
    "STD128_MKNTRU": BinFHEContextParams(2, 27, 4096, 765, 45181, 45181, STD_NTRU, 32, 1 << 7, 32),
    "STD128_MKNTRU_2": BinFHEContextParams(4, 27, 4096, 765, 45181, 45181, STD_NTRU, 32, 1 << 7, 32),
    "STD128_MKNTRU_3": BinFHEContextParams(8, 27, 4096, 765, 45181, 45181, STD_NTRU, 32, 1 << 6, 32),
    "STD128_MKNTRU_4": BinFHEContextParams(16, 27, 4096, 765, 45181, 45181, STD_NTRU, 32, 1 << 5, 32),
    "STD128_MKNTRU_LWE": BinFHEContextParams(2, 27, 4096, 635, 32749, 32749, STD_DEV, 32, 1 << 9, 2),
    "STD128_MKNTRU_LWE_2": BinFHEContextParams(4, 27, 4096, 635, 32749, 32749, STD_DEV, 32, 1 << 9, 2),
    "STD128_MKNTRU_LWE_3": BinFHEContextParams(8, 27, 4096, 635, 32749, 32749, STD_DEV, 32, 1 << 9, 2),
    "STD128_MKNTRU_LWE_4": BinFHEContextParams(16, 27, 4096, 635, 32749, 32749, STD_DEV, 32, 1 << 7, 2),
    "STD100_MKNTRU": BinFHEContextParams(2, 27, 4096, 560, 45181, 45181, STD_NTRU2, 32, 1 << 9, 32),
    "STD100_MKNTRU_2": BinFHEContextParams(4, 27, 4096, 560, 45181, 45181, STD_NTRU2, 32, 1 << 9, 32),
    "STD100_MKNTRU_3": BinFHEContextParams(8, 27, 4096, 560, 45181, 45181, STD_NTRU2, 32, 1 << 9, 32),
    "STD100_MKNTRU_4": BinFHEContextParams(16, 27, 4096, 560, 45181, 45181, STD_NTRU2, 32, 1 << 9, 32),
    "STD100_MKNTRU_LWE": BinFHEContextParams(2, 27, 4096, 500, 32749, 32749, STD_DEV, 32, 1 << 9, 2),
    "STD100_MKNTRU_LWE_2": BinFHEContextParams(4, 27, 4096, 500, 32749, 32749, STD_DEV, 32, 1 << 9, 2),
    "STD100_MKNTRU_LWE_3": BinFHEContextParams(8, 27, 4096, 500, 32749, 32749, STD_DEV, 32, 1 << 9, 2),
    "STD100_MKNTRU_LWE_4": BinFHEContextParams(16, 27, 4096, 500, 32749, 32749, STD_DEV, 32, 1 << 9, 2),
}


# ---- parameter derivation (host logic, integers only) ----
def _is_prime(n: int) -> bool:
    if n < 2:
        return False
    for p in (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37):
        if n % p == 0:
            return n == p
    d, s = n - 1, 0
    while d % 2 == 0:
        d //= 2
        s += 1
    for a in (2, 3, 5, 7, 11, 13):
        x = pow(a, d, n)
        if x in (1, n - 1):
            continue
        for _ in range(s - 1):
            x = x * x % n
            if x == n - 1:
                break
        else:
            return False
    return True


def first_prime(nbits: int, m: int) -> int:
    """FirstPrime (src/core/include/math/nbtheory-impl.h:334-358)."""
    q = 1 << nbits
    r = q % m
    q2 = q + 1 + ((m - r) if r else 0)
    while not _is_prime(q2):
        q2 += m
    return q2


def previous_prime(q: int, m: int) -> int:
    """PreviousPrime (nbtheory-impl.h:371-379)."""
    c = q - m
    while not _is_prime(c):
        c -= m
    return c


def root_of_unity(m: int, Q: int) -> int:
    """RootOfUnity (nbtheory-impl.h:183-231): the minimal primitive m-th root (m a power of two)."""
    g = next(c for c in (pow(x, (Q - 1) // m, Q) for x in range(2, Q)) if pow(c, m // 2, Q) == Q - 1)
    best, x, g2 = g, g, g * g % Q
    for _ in range(1, m, 2):
        best = min(best, x)
        x = x * g2 % Q
    return best


def derive_params(row: BinFHEContextParams, method: str, device: int = -1) -> Params:
    if method not in (MKNTRU, MKNTRU_B, MKNTRU_LWE):
        raise config_error("method is invalid")  # binfhe-base-scheme.h:137-151
    N = row.cyclOrder // 2
    Q = previous_prime(first_prime(row.numberBits, row.cyclOrder), row.cyclOrder)  # binfhecontext.cpp:157-158
    psi = root_of_unity(row.cyclOrder, Q)
    if row.gadgetBase & (row.gadgetBase - 1):
        raise config_error("Gadget base should be a power of two.")  # mk-cryptoparameters.h:138-139
    digitsG = int(ceil(log(float(Q)) / log(float(row.gadgetBase))))  # mk-cryptoparameters.h:141-142
    dKS = int(ceil(log(float(row.modKS)) / log(float(row.baseKS))))  # mntru-pke.cpp:771
    return Params(k=row.numUser, N=N, Q=Q, psi=psi, q=row.mod, qKS=row.modKS, n=row.latticeParam,
                  baseG_log2=row.gadgetBase.bit_length() - 1, digitsG=digitsG, baseKS=row.baseKS, dKS=dKS,
                  variant=_lib.MKFHE_VARIANT_XZW if method == MKNTRU else _lib.MKFHE_VARIANT_XZW_B,
                  scheme=_lib.MKFHE_SCHEME_MKLWE if method == MKNTRU_LWE else _lib.MKFHE_SCHEME_MNTRU,
                  device=device)


def _ptr(a):
    """host numpy array or torch tensor (host or CUDA) -> (void*, words)"""
    if a is None:
        return None, 0
    if isinstance(a, np.ndarray):
        if a.dtype != np.uint32 or not a.flags["C_CONTIGUOUS"]:
            raise config_error("expected a C-contiguous uint32 array")
        return C.c_void_p(a.ctypes.data), a.size
    # torch tensor: int32/uint32 storage, contiguous
    if not a.is_contiguous() or a.element_size() != 4:
        raise config_error("expected a contiguous 32-bit tensor")
    return C.c_void_p(a.data_ptr()), a.numel()


class BinFHEContext:
    """GPU context for one parameter set on one device (one per GPU; calls are serialised by the caller)."""

    def __init__(self):
        self._L = None
        self._h = None
        self.params: Params | None = None

    # -- BinFHEContext::GenerateBinFHEContext(BINFHE_PARAMSET, BINFHE_METHOD), binfhecontext.cpp:83-194
    def GenerateBinFHEContext(self, paramset, method: str, device: int = -1):
        row = PARAMSETS.get(paramset) if isinstance(paramset, str) else paramset
        if row is None:
            raise config_error(f"ERROR: Unknown parameter set [{paramset}] for FHEW.")  # binfhecontext.cpp:149-152
        return self.GenerateFromParams(derive_params(row, method, device))

    def GenerateFromParams(self, p: Params):
        self.close()
        self._L = _lib.load()
        h = C.c_void_p()
        rc = self._L.mkfhe_ctx_create(C.byref(p), C.byref(h))
        if rc != 0:
            raise MkfheError(rc, "mkfhe_ctx_create", self._L.mkfhe_last_error(None).decode())
        self._h, self.params = h, p
        return self

    def close(self):
        if self._h:
            self._L.mkfhe_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, what):
        if rc != 0:
            raise MkfheError(rc, what, self._L.mkfhe_last_error(self._h).decode())

    @property
    def is_lwe(self) -> bool:
        return self.params.scheme == _lib.MKFHE_SCHEME_MKLWE

    # -- key upload: the flattened bytes of what BinFHEContext::MKBTKeyGen / ctGateGen produce in the reference
    def LoadKeys(self, bk, pkey, ksk=None, kskA=None, kskB=None, ctnand=None):
        p, n = _ptr(bk)
        self._ck(self._L.mkfhe_load_bk(self._h, p, n), "mkfhe_load_bk")
        p, n = _ptr(pkey)
        self._ck(self._L.mkfhe_load_pkey(self._h, p, n), "mkfhe_load_pkey")
        if ksk is not None:
            p, n = _ptr(ksk)
            self._ck(self._L.mkfhe_load_ksk_mntru(self._h, p, n), "mkfhe_load_ksk_mntru")
        if kskA is not None:
            pa, na = _ptr(kskA)
            pb, nb = _ptr(kskB)
            self._ck(self._L.mkfhe_load_ksk_mklwe(self._h, pa, na, pb, nb), "mkfhe_load_ksk_mklwe")
        if ctnand is not None:
            p, n = _ptr(ctnand)
            self._ck(self._L.mkfhe_load_ctnand(self._h, p, n), "mkfhe_load_ctnand")

    # -- BinFHEContext::EvalBinGate(gate, ct1, ct2), binfhecontext.cpp:415-426, batched over the leading axis
    def EvalBinGate(self, gate, ct1, ct2, b1=None, b2=None):
        """ct1, ct2: uint32 [B, k, n] (or [k, n]); b1, b2: uint32 [B] for MKNTRU_LWE.  Returns out (and bout)."""
        if gate != NAND:
            raise config_error("Support NAND gate Only")  # binfhe-base-scheme.cpp:341-342
        if ct1 is ct2:
            raise config_error("Input ciphertexts should be independant")  # binfhe-base-scheme.cpp:382-383,471-472
        P = self.params
        c1 = np.ascontiguousarray(ct1, dtype=np.uint32).reshape(-1, P.k, P.n)
        c2 = np.ascontiguousarray(ct2, dtype=np.uint32).reshape(-1, P.k, P.n)
        if c1.shape != c2.shape:
            raise config_error("ciphertext shapes differ")
        B = c1.shape[0]
        out = np.empty_like(c1)
        if self.is_lwe:
            bb1 = np.ascontiguousarray(b1, dtype=np.uint32).reshape(B)
            bb2 = np.ascontiguousarray(b2, dtype=np.uint32).reshape(B)
            bout = np.empty(B, dtype=np.uint32)
            self._ck(self._L.mkfhe_eval_nand(self._h, _ptr(c1)[0], _ptr(bb1)[0], _ptr(c2)[0], _ptr(bb2)[0], _ptr(out)[0],
                                             _ptr(bout)[0], B), "mkfhe_eval_nand")
            return out.reshape(np.shape(ct1)), (bout if np.ndim(b1) else int(bout[0]))
        self._ck(self._L.mkfhe_eval_nand(self._h, _ptr(c1)[0], None, _ptr(c2)[0], None, _ptr(out)[0], None, B),
                 "mkfhe_eval_nand")
        return out.reshape(np.shape(ct1))

    def EvalBinGateDevice(self, ct1, ct2, out, b1=None, b2=None, bout=None):
        """Same on CUDA tensors (int32 storage, [B, k, n]); asynchronous on the context's stream."""
        B = ct1.numel() // (self.params.k * self.params.n)
        g = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        self._ck(self._L.mkfhe_eval_nand_dev(self._h, g(ct1), g(b1), g(ct2), g(b2), g(out), g(bout), B), "mkfhe_eval_nand_dev")

    # -- circuits of NAND gates: what a caller of the reference writes as a loop over EvalBinGate (binfhecontext.cpp:415-419)
    def EvalCircuit(self, wires, gates, b=None):
        """wires: uint32 [W, k, n] with the primary inputs filled in; gates: int [G, 3] rows (in1, in2, out) in topological order;
        b: uint32 [W] for MKNTRU_LWE.  Gate g sets wires[out] = NAND(wires[in1], wires[in2]); every level of the netlist runs as
        one batch.  Returns (wires, b, levels) -- new arrays, the inputs are not modified."""
        P = self.params
        w = np.ascontiguousarray(wires, dtype=np.uint32).reshape(-1, P.k, P.n).copy()
        g = np.ascontiguousarray(gates, dtype=np.uint32).reshape(-1, 3)
        i1, i2, o = (np.ascontiguousarray(g[:, j]) for j in range(3))
        if len(g) and bool((i1 == i2).any()):
            raise config_error("Input ciphertexts should be independant")  # binfhe-base-scheme.cpp:382-383,471-472
        bb = None
        if self.is_lwe:
            if b is None:
                raise config_error("MKLWE ciphertexts carry b")
            bb = np.ascontiguousarray(b, dtype=np.uint32).reshape(w.shape[0]).copy()
        lv = C.c_uint32(0)
        self._ck(self._L.mkfhe_eval_circuit(self._h, _ptr(w)[0], _ptr(bb)[0], w.shape[0], _ptr(i1)[0], _ptr(i2)[0], _ptr(o)[0], len(g),
                                            C.byref(lv)), "mkfhe_eval_circuit")
        return w, bb, int(lv.value)

    def EvalCircuitDevice(self, wires, gates, b=None) -> int:
        """Same on CUDA tensors (int32 storage, [W, k, n] and [W]), in place and asynchronous; gates stays a host array."""
        g = np.ascontiguousarray(gates, dtype=np.uint32).reshape(-1, 3)
        i1, i2, o = (np.ascontiguousarray(g[:, j]) for j in range(3))
        W = wires.numel() // (self.params.k * self.params.n)
        lv = C.c_uint32(0)
        dp = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        self._ck(self._L.mkfhe_eval_circuit_dev(self._h, dp(wires), dp(b), W, _ptr(i1)[0], _ptr(i2)[0], _ptr(o)[0], len(g), C.byref(lv)),
                 "mkfhe_eval_circuit_dev")
        return int(lv.value)

    def EvalCircuitDeviceStatic(self, wires, gates, b=None) -> int:
        """EvalCircuitDevice without device allocation (mkfhe_eval_circuit_dev_static: the scratch of the level batches lives in
        the library's file-scope device array; no counterpart in the reference).  Raises MkfheError (MKFHE_EINVAL) for a circuit
        beyond the array's size; calls that overlap on one GPU race on the shared array."""
        g = np.ascontiguousarray(gates, dtype=np.uint32).reshape(-1, 3)
        i1, i2, o = (np.ascontiguousarray(g[:, j]) for j in range(3))
        W = wires.numel() // (self.params.k * self.params.n)
        lv = C.c_uint32(0)
        dp = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        self._ck(self._L.mkfhe_eval_circuit_dev_static(self._h, dp(wires), dp(b), W, _ptr(i1)[0], _ptr(i2)[0], _ptr(o)[0], len(g),
                                                       C.byref(lv)), "mkfhe_eval_circuit_dev_static")
        return int(lv.value)

    # -- BinFHEContext::Decrypt (binfhecontext.h:318-338 -> mntru-pke.cpp:306-341 / mklwe-pke.cpp:66-97), batched
    def LoadSecretKey(self, sk):
        """sk: uint32 [k, n] -- MNTRUPrivateKey::GetF_col0() or MKLWEPrivateKey::GetElement() of the decrypting parties"""
        p, n = _ptr(sk)
        self._ck(self._L.mkfhe_load_sk(self._h, p, n), "mkfhe_load_sk")

    def Decrypt(self, ct, b=None, p: int = 4):
        """ct: uint32 [B, k, n] (or [k, n]); b: uint32 [B] for MKNTRU_LWE.  Returns the plaintexts, uint32 [B]."""
        P = self.params
        c = np.ascontiguousarray(ct, dtype=np.uint32).reshape(-1, P.k, P.n)
        B = c.shape[0]
        m = np.empty(B, dtype=np.uint32)
        bb = None
        if self.is_lwe:
            if b is None:
                raise config_error("MKLWE ciphertexts carry b")
            bb = np.ascontiguousarray(b, dtype=np.uint32).reshape(B)
        self._ck(self._L.mkfhe_decrypt(self._h, _ptr(c)[0], _ptr(bb)[0], int(p), _ptr(m)[0], B), "mkfhe_decrypt")
        return m

    def DecryptDevice(self, ct, m, b=None, p: int = 4):
        B = ct.numel() // (self.params.k * self.params.n)
        g = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        self._ck(self._L.mkfhe_decrypt_dev(self._h, g(ct), g(b), int(p), g(m), B), "mkfhe_decrypt_dev")

    # -- UniEncAccumulator::EvalAcc (mk-acc.h:67-71), batched
    def EvalAcc(self, acc, rot, rot_mod: int):
        P = self.params
        a = np.ascontiguousarray(acc, dtype=np.uint32).reshape(-1, P.k, P.N)
        r = np.ascontiguousarray(rot, dtype=np.uint32).reshape(-1, P.k, P.n)
        out = np.empty_like(a)
        self._ck(self._L.mkfhe_eval_acc(self._h, _ptr(a)[0], _ptr(r)[0], int(rot_mod), _ptr(out)[0], a.shape[0]), "mkfhe_eval_acc")
        return out.reshape(np.shape(acc))

    # -- NativePoly::SetFormat, batched
    def NTT(self, polys, inverse: bool = False):
        a = np.ascontiguousarray(polys, dtype=np.uint32).copy()
        self._ck(self._L.mkfhe_ntt(self._h, _ptr(a)[0], a.size // self.params.N, 1 if inverse else 0), "mkfhe_ntt")
        return a

    def NTTDevice(self, polys, inverse: bool = False):
        self._ck(self._L.mkfhe_ntt_dev(self._h, C.c_void_p(polys.data_ptr()), polys.numel() // self.params.N,
                                       1 if inverse else 0), "mkfhe_ntt_dev")

    def SignedDigitDecompose(self, polys):
        a = np.ascontiguousarray(polys, dtype=np.uint32).reshape(-1, self.params.N)
        out = np.empty((a.shape[0], self.params.digitsG - 1, self.params.N), dtype=np.uint32)
        self._ck(self._L.mkfhe_decompose(self._h, _ptr(a)[0], _ptr(out)[0], a.shape[0]), "mkfhe_decompose")
        return out

    # -- resident key images (include/mkfhe_b200.h "Resident key images"): what mkfhe_b200/shard.py broadcasts between GPUs
    def KeyImages(self):
        """[(which, device pointer, bytes)] of this context's key images, allocating those not loaded yet."""
        P = self.params
        which = [_lib.MKFHE_IMAGE_BK, _lib.MKFHE_IMAGE_PKEY, _lib.MKFHE_IMAGE_KSK]
        if P.scheme == _lib.MKFHE_SCHEME_MNTRU:
            which.append(_lib.MKFHE_IMAGE_CTNAND)
        out = []
        for w in which:
            ptr, nb = C.c_void_p(), C.c_uint64()
            self._ck(self._L.mkfhe_key_image(self._h, w, C.byref(ptr), C.byref(nb)), "mkfhe_key_image")
            out.append((w, int(ptr.value), int(nb.value)))
        return out

    def CommitKeyImages(self, which=None):
        for w in (which if which is not None else [x[0] for x in self.KeyImages()]):
            self._ck(self._L.mkfhe_key_image_commit(self._h, int(w)), "mkfhe_key_image_commit")

    # -- plumbing
    def set_stream(self, cuda_stream_handle: int | None):
        """All later work of this context goes to `cuda_stream_handle` (a cudaStream_t as an integer, e.g.
        torch.cuda.Stream().cuda_stream).  None / 0: the context's own stream, an ordinary blocking stream and therefore ordered
        with the legacy default stream (torch's default stream); lib.CUDA_STREAM_LEGACY (1) selects the legacy default stream
        itself.  *Device methods read inputs and write outputs in stream order on that stream and do not wait."""
        self._ck(self._L.mkfhe_ctx_set_stream(self._h, C.c_void_p(cuda_stream_handle) if cuda_stream_handle else None),
                 "mkfhe_ctx_set_stream")

    def synchronize(self):
        self._ck(self._L.mkfhe_ctx_synchronize(self._h), "mkfhe_ctx_synchronize")

    def last_gate_timing(self):
        """(blind_rotation_ms, key_switch_ms) of the last gate batch, from CUDA events on the context's stream"""
        a, b = C.c_float(0), C.c_float(0)
        self._ck(self._L.mkfhe_last_gate_timing(self._h, C.byref(a), C.byref(b)), "mkfhe_last_gate_timing")
        return float(a.value), float(b.value)

    def last_gate_sms(self) -> int:
        """SMs per gate of the last blind-rotation launch: 1, or the thread-block cluster size (2 / 4 / 8) of a small batch"""
        return int(self._L.mkfhe_last_gate_sms(self._h))

    def launch_count(self) -> int:
        return int(self._L.mkfhe_launch_count(self._h))

    def device_bytes(self) -> int:
        return int(self._L.mkfhe_device_bytes(self._h))

    def measure_int32_peak(self) -> float:
        v = C.c_double(0)
        self._ck(self._L.mkfhe_measure_int32_peak(self._h, C.byref(v)), "mkfhe_measure_int32_peak")
        return float(v.value)

    def measure_butterfly_peak(self) -> float:
        """Lazy Shoup butterflies per second this GPU sustains on registers alone (mkfhe_measure_butterfly_peak)."""
        v = C.c_double(0)
        self._ck(self._L.mkfhe_measure_butterfly_peak(self._h, C.byref(v)), "mkfhe_measure_butterfly_peak")
        return float(v.value)

    # -- key generation with caller-supplied randomness (UniEncAccumulatorXZW{,_B}::KeyGenAcc, mk-acc-xzw.cpp:38-87,132-228;
    #    Pkey: binfhe-base-scheme.cpp:250-269).  Host numpy arrays in, host numpy arrays out; load=True also installs the keys.
    def KeyGenAcc(self, sinv, crs, r, e, m, load: bool = False):
        P = self.params
        ell = P.digitsG - 1
        S = 2 if P.variant == 0 else 1
        E = P.k * S * P.n + 1
        u32 = lambda a, shape: np.ascontiguousarray(a, dtype=np.uint32).reshape(shape)
        sinv, crs = u32(sinv, (P.k, P.N)), u32(crs, (ell, P.N))
        r, e = u32(r, (E, P.N)), u32(e, (E, ell, 2, P.N))
        m = np.ascontiguousarray(m, dtype=np.uint8).reshape(E)
        out = np.empty((E, ell, 2, P.N), dtype=np.uint32)
        self._ck(self._L.mkfhe_keygen_bk(self._h, _ptr(sinv)[0], _ptr(crs)[0], _ptr(r)[0], _ptr(e)[0],
                                         C.c_void_p(m.ctypes.data), _ptr(out)[0], int(load)), "mkfhe_keygen_bk")
        return out.reshape(-1)

    def KeyGenPkey(self, s, crs, e, load: bool = False):
        P = self.params
        ell = P.digitsG - 1
        u32 = lambda a, shape: np.ascontiguousarray(a, dtype=np.uint32).reshape(shape)
        s, crs, e = u32(s, (P.k, P.N)), u32(crs, (ell, P.N)), u32(e, (P.k, ell, P.N))
        out = np.empty((P.k, ell, P.N), dtype=np.uint32)
        self._ck(self._L.mkfhe_keygen_pkey(self._h, _ptr(s)[0], _ptr(crs)[0], _ptr(e)[0], _ptr(out)[0], int(load)),
                 "mkfhe_keygen_pkey")
        return out.reshape(-1)

    # -- key-switch key generation with caller-supplied noise (mntru-pke.cpp:624-760 digit-1 slice; mklwe-pke.cpp:176-256)
    def KeySwitchGen2(self, s_q, e, finv, load: bool = False):
        P = self.params
        L = P.N * P.dKS
        u32 = lambda a, shape: np.ascontiguousarray(a, dtype=np.uint32).reshape(shape)
        s_q, e, finv = u32(s_q, (P.k, P.N)), u32(e, (P.k, L, P.n)), u32(finv, (P.k, P.n, P.n))
        out = np.empty((P.k, L, P.n), dtype=np.uint32)
        self._ck(self._L.mkfhe_keygen_ksk_mntru(self._h, _ptr(s_q)[0], _ptr(e)[0], _ptr(finv)[0], _ptr(out)[0], int(load)),
                 "mkfhe_keygen_ksk_mntru")
        return out.reshape(-1)

    def KeySwitchGenLWE(self, A, e, s, sN_q, load: bool = False):
        P = self.params
        rows = P.k * P.N * P.baseKS * P.dKS
        u32 = lambda a, shape: np.ascontiguousarray(a, dtype=np.uint32).reshape(shape)
        A, e, s, sN_q = u32(A, (rows, P.n)), u32(e, (rows,)), u32(s, (P.k, P.n)), u32(sN_q, (P.k, P.N))
        out = np.empty(rows, dtype=np.uint32)
        self._ck(self._L.mkfhe_keygen_ksk_mklwe(self._h, _ptr(A)[0], _ptr(e)[0], _ptr(s)[0], _ptr(sN_q)[0], _ptr(out)[0],
                                                int(load)), "mkfhe_keygen_ksk_mklwe")
        return out

    # -- binary key file (mkfhe_b200/csrc/keyfile.h): the flattened layouts, checksummed; replaces the reference's cereal
    #    serialisation of the key object graphs for the GPU path
    def LoadKeyFile(self, path: str):
        self._ck(self._L.mkfhe_keyfile_load(self._h, os.fsencode(path)), "mkfhe_keyfile_load")


def SaveKeyFile(path: str, params: Params, bk, pkey, ksk=None, kskA=None, kskB=None, ctnand=None):
    """Write a key file for `params` (no GPU needed).  MNTRU: ksk (digit-1 slice) + ctnand; MKNTRU_LWE: kskA + kskB."""
    L = _lib.load()
    a = kskA if kskA is not None else ksk
    args = []
    for x in (bk, pkey, a, kskB, ctnand):
        p, n = _ptr(None if x is None else np.ascontiguousarray(x, dtype=np.uint32))
        args += [p, n]
    if L.mkfhe_keyfile_write(os.fsencode(path), C.byref(params), *args) != 0:
        raise _lib.MkfheError(-1, "mkfhe_keyfile_write", f"cannot write {path}")


def KeyFileInfo(path: str):
    """(Params, {section name: words}) of a key file, verifying every checksum (no GPU needed)."""
    L = _lib.load()
    p = Params()
    w = (C.c_uint64 * 7)()
    if L.mkfhe_keyfile_info(os.fsencode(path), C.byref(p), w) != 0:
        raise _lib.MkfheError(-1, "mkfhe_keyfile_info", f"{path}: unreadable, not a key file, or checksum mismatch")
    names = {1: "bk", 2: "pkey", 3: "ksk", 4: "kskA", 5: "kskB", 6: "ctnand"}
    return p, {names[i]: int(w[i]) for i in names if w[i]}


def SaveCiphertexts(path: str, params: Params, ct, b=None):
    """Write ciphertexts ct [n, k, n_lwe] (and b [n] for MKNTRU_LWE) as one checksummed wire-table file (no GPU needed)."""
    L = _lib.load()
    c = np.ascontiguousarray(ct, dtype=np.uint32).reshape(-1, params.k, params.n)
    bb = None if b is None else np.ascontiguousarray(b, dtype=np.uint32).reshape(c.shape[0])
    if L.mkfhe_ctfile_write(os.fsencode(path), C.byref(params), _ptr(c)[0], _ptr(bb)[0], c.shape[0]) != 0:
        raise _lib.MkfheError(-1, "mkfhe_ctfile_write", f"cannot write {path} (MKNTRU_LWE ciphertexts need b)")


def LoadCiphertexts(path: str, params: Params | None = None):
    """(Params, ct [n, k, n_lwe], b [n] or None) of a wire-table file; `params`, when given, must match the file's."""
    L = _lib.load()
    p, n = Params(), C.c_uint64()
    if L.mkfhe_ctfile_info(os.fsencode(path), C.byref(p), C.byref(n)) != 0:
        raise _lib.MkfheError(-1, "mkfhe_ctfile_info", f"{path}: unreadable, not a ciphertext file, or checksum mismatch")
    ct = np.empty((n.value, p.k, p.n), dtype=np.uint32)
    b = np.empty(n.value, dtype=np.uint32) if p.scheme == _lib.MKFHE_SCHEME_MKLWE else None
    got = C.c_uint64()
    if L.mkfhe_ctfile_read(os.fsencode(path), C.byref(params) if params is not None else None, _ptr(ct)[0], _ptr(b)[0], n.value,
                           C.byref(got)) != 0:
        raise _lib.MkfheError(-1, "mkfhe_ctfile_read", f"{path}: written for a different parameter set")
    return p, ct, b


def CircuitLevels(n_wires: int, gates):
    """Levelisation EvalCircuit applies (no GPU needed): gates int [G, 3] rows (in1, in2, out) -> (levels, level of each gate).
    Raises config_error for a netlist EvalCircuit refuses (index out of range, in1 == in2, a wire written twice or read
    before the gate that writes it)."""
    L = _lib.load()
    g = np.ascontiguousarray(gates, dtype=np.uint32).reshape(-1, 3)
    i1, i2, o = (np.ascontiguousarray(g[:, j]) for j in range(3))
    lev = np.zeros(len(g), dtype=np.uint32)
    n = L.mkfhe_circuit_levels(int(n_wires), _ptr(i1)[0], _ptr(i2)[0], _ptr(o)[0], len(g), _ptr(lev)[0])
    if n < 0:
        raise config_error("not a levelisable NAND netlist")
    return int(n), lev


def _encrypt_methods():
    # -- batched Encrypt with caller-supplied noise (mntru-pke.cpp:158-205, mklwe-pke.cpp:36-64)
    def EncryptBatch(self, m, noise, secret, a=None, pmod: int = 4):
        """MNTRU: noise = e [B][k][n], secret = F^-1 [k][n][n] -> ct [B][k][n].
        MKNTRU_LWE: a [B][k][n] (the sampled vectors = the ciphertext's a part), noise = err [B], secret = s [k][n] -> b [B]."""
        P = self.params
        m = np.ascontiguousarray(m, dtype=np.uint8).reshape(-1)
        B = m.size
        u32 = lambda x, shape: np.ascontiguousarray(x, dtype=np.uint32).reshape(shape)
        mp = C.c_void_p(m.ctypes.data)
        if not self.is_lwe:
            e, finv = u32(noise, (B, P.k, P.n)), u32(secret, (P.k, P.n, P.n))
            out = np.empty((B, P.k, P.n), dtype=np.uint32)
            self._ck(self._L.mkfhe_encrypt_mntru(self._h, _ptr(finv)[0], _ptr(e)[0], mp, pmod, _ptr(out)[0], B), "mkfhe_encrypt_mntru")
            return out
        a, err, s = u32(a, (B, P.k, P.n)), u32(noise, (B,)), u32(secret, (P.k, P.n))
        out = np.empty(B, dtype=np.uint32)
        self._ck(self._L.mkfhe_encrypt_mklwe(self._h, _ptr(s)[0], _ptr(a)[0], _ptr(err)[0], mp, pmod, _ptr(out)[0], B),
                 "mkfhe_encrypt_mklwe")
        return out

    BinFHEContext.EncryptBatch = EncryptBatch


_encrypt_methods()
