"""CPU: the C-ABI library loads and exports exactly the symbols include/mkfhe_b200.h declares; without a GPU it refuses to
create a context (no CPU fallback); host-side parameter derivation matches the reference's constants."""
import ctypes
import os
import re

import numpy as np
import pytest

import mkfhe_b200 as mk
from mkfhe_b200 import lib as mklib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "mkfhe_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mkfhe_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported():
    declared = _declared()
    assert declared == sorted(mklib.SYMBOLS)
    L = ctypes.CDLL(mklib.LIB_PATH)
    for s in declared:
        assert hasattr(L, s), s
    assert b"sm_100a" in mklib.load().mkfhe_version()


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(mk.MkfheError) as e:
        mk.BinFHEContext().GenerateBinFHEContext("STD100_MKNTRU", mk.MKNTRU)
    assert e.value.code == -2  # MKFHE_ENODEV


def test_parameter_derivation_matches_reference_constants():
    # SURVEY section 8 (probe-verified against the reference): Q(27,4096)=134176769, psi=100530; Q(27,2048)=134215681, psi=282116
    p = mk.derive_params(mk.PARAMSETS["STD100_MKNTRU"], mk.MKNTRU)
    assert (p.Q, p.psi, p.digitsG, p.dKS, p.N, p.n, p.k) == (134176769, 100530, 3, 4, 2048, 560, 2)
    p = mk.derive_params(mk.PARAMSETS["STD100_MKNTRU_LWE"], mk.MKNTRU_LWE)
    assert (p.Q, p.psi, p.digitsG, p.dKS, p.variant, p.scheme) == (134176769, 100530, 3, 3, 1, 1)
    from mkfhe_b200.context import first_prime, previous_prime, root_of_unity

    Q = previous_prime(first_prime(27, 2048), 2048)
    assert (Q, root_of_unity(2048, Q)) == (134215681, 282116)
    assert mk.derive_params(mk.PARAMSETS["STD128_MKNTRU_4"], mk.MKNTRU).digitsG == 6
    with pytest.raises(mk.config_error):
        mk.derive_params(mk.PARAMSETS["STD100_MKNTRU"], "GINX")


def test_keyfile_write_info_roundtrip_and_corruption(tmp_path):
    """The binary key file needs no GPU to write or to verify: sizes and parameters come back, a flipped
    payload byte is caught by the section checksum, a foreign file is rejected."""
    import numpy as np

    import mkfhe_b200 as mk
    from util import load_golden

    g = load_golden("toy_mkntru_k4_N256")
    p = g["p"]
    P = mk.Params(k=p.k, N=p.N, Q=p.Q, psi=p.psi, q=p.q, qKS=p.qKS, n=p.n, baseG_log2=p.baseG_log2, digitsG=p.digitsG,
                  baseKS=p.baseKS, dKS=p.dKS, variant=p.variant, scheme=p.scheme, device=-1)
    path = str(tmp_path / "keys.mkfhe")
    mk.SaveKeyFile(path, P, g["bk"], g["pkey"], ksk=g["ksk"], ctnand=g["ctnand"])
    Pr, sizes = mk.KeyFileInfo(path)
    assert (Pr.k, Pr.N, Pr.Q, Pr.n, Pr.scheme, Pr.variant) == (p.k, p.N, p.Q, p.n, p.scheme, p.variant)
    assert sizes == {"bk": np.size(g["bk"]), "pkey": np.size(g["pkey"]), "ksk": np.size(g["ksk"]), "ctnand": np.size(g["ctnand"])}
    raw = bytearray(open(path, "rb").read())
    raw[len(raw) // 2] ^= 1
    open(path, "wb").write(raw)
    with pytest.raises(mk.MkfheError):
        mk.KeyFileInfo(path)
    open(path, "wb").write(b"not a key file at all, just some bytes" * 4)
    with pytest.raises(mk.MkfheError):
        mk.KeyFileInfo(path)


def test_circuit_levelisation_needs_no_gpu():
    """mkfhe_circuit_levels: the host-side levelisation behind mkfhe_eval_circuit (include/mkfhe_b200.h)."""
    import mkfhe_b200 as mk

    xor = [(0, 1, 2), (0, 2, 3), (1, 2, 4), (3, 4, 5)]
    n, lev = mk.CircuitLevels(6, xor)
    assert n == 3 and list(lev) == [0, 1, 1, 2]
    # two independent chains interleave into shared levels
    n, lev = mk.CircuitLevels(8, [(0, 1, 4), (2, 3, 5), (4, 5, 6), (0, 6, 7)])
    assert n == 3 and list(lev) == [0, 0, 1, 2]
    assert mk.CircuitLevels(3, np.zeros((0, 3), dtype=np.uint32))[0] == 0
    for bad in ([(0, 0, 2)],              # one ciphertext on both inputs (binfhe-base-scheme.cpp:382-383)
                [(0, 1, 6)],              # index out of range
                [(0, 1, 2), (0, 1, 2)],   # a wire written twice
                [(0, 2, 3), (0, 1, 2)],   # read before written
                [(0, 1, 0)]):             # a gate overwriting its own input
        with pytest.raises(mk.config_error):
            mk.CircuitLevels(6, bad)


def test_header_is_plain_c(tmp_path):
    """include/mkfhe_b200.h is the drop-in boundary: it must compile as C (no C++ types, no torch types) and as C++."""
    import subprocess

    src = tmp_path / "use.c"
    src.write_text('#include "mkfhe_b200.h"\nint main(void) { mkfhe_params p; (void)p; return (int)(MKFHE_STATIC_SCRATCH_WORDS == 0); }\n')
    inc = os.path.join(ROOT, "include")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-fsyntax-only", "-I", inc, str(src)])
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-fsyntax-only", "-x", "c++", "-I", inc, str(src)])


def test_static_scratch_sizing_needs_no_gpu():
    """mkfhe_circuit_static_words / mkfhe_static_scratch_words (the allocation-free circuit entry point): host arithmetic only"""
    L = mklib.load()
    cap = L.mkfhe_static_scratch_words()
    assert cap == 2 << 20
    P = mk.derive_params(mk.PARAMSETS["STD100_MKNTRU_LWE"], mk.MKNTRU_LWE)
    pad = lambda w: (w + 63) & ~63
    for gates, widest in ((1, 1), (100, 7), (3000, 600)):
        want = pad(3 * gates) + 3 * pad(widest * P.k * P.n) + 3 * pad(widest)
        assert L.mkfhe_circuit_static_words(ctypes.byref(P), gates, widest) == want
    Pn = mk.derive_params(mk.PARAMSETS["STD100_MKNTRU_4"], mk.MKNTRU)  # MNTRU: no b column
    assert L.mkfhe_circuit_static_words(ctypes.byref(Pn), 10, 10) == pad(30) + 3 * pad(10 * Pn.k * Pn.n)
    assert L.mkfhe_circuit_static_words(ctypes.byref(P), 4 * 695, 695) <= cap < L.mkfhe_circuit_static_words(ctypes.byref(P), 4 * 696, 696)
