"""CPU: the plain-C oracle (oracle/mkfhe_oracle.c) against
  (1) the committed golden fixtures, which were produced by the compiled reference (tests/golden/make_golden.py),
  (2) the reference's own known-answer test for the transform (src/core/unittest/UnitTestTransform.cpp:59-94),
  (3) the compiled reference itself (oracle/_ref) when it is present, on fresh random keys and inputs.
This is what pins the oracle; the GPU suite then compares the CUDA path against the oracle.
"""
import numpy as np
import pytest

from oracle import portapi, refapi
from util import golden_names, load_golden, make_oracle, params_from_ref


@pytest.mark.parametrize("name", golden_names())
def test_oracle_matches_golden(name):
    g = load_golden(name)
    p = g["p"]
    o = make_oracle(p)
    assert portapi.derive_modulus(p.numberBits, p.cyclOrder) == (p.Q, p.psi)
    assert (o.ntt(g["ntt_in"], False) == g["ntt_fwd"]).all()
    assert (o.ntt(g["ntt_in"], True) == g["ntt_inv"]).all()
    for x, y in zip(g["dec_in"], g["dec_out"]):
        assert (o.decompose(x) == y).all()
    for a, r, e in zip(g["acc_in"], g["acc_rot"], g["acc_out"]):
        assert (o.eval_acc(g["bk"], g["pkey"], a, r, p.rot_mod) == e).all()
    for i in range(len(g["g_out"])):
        out, bout = o.eval_nand(g["bk"], g["pkey"], g["kskA_any"], g["kskB_any"], g["ctnand_any"], g["g_ct1"][i],
                                g["g_b1"][i], g["g_ct2"][i], g["g_b2"][i])
        assert (out == g["g_out"][i]).all() and bout == int(g["g_bout"][i])
        assert o.decrypt(g["sk"], out, bout) == int(g["g_bits"][i])
        m0, m1 = (int(x) for x in g["g_m"][i])
        assert int(g["g_bits"][i]) == 1 - (m0 & m1)  # the reference's own result is NAND


def test_reference_transform_known_answer():
    # UnitTestTransform.cpp:59-94: (1 + 2x + 4x^2 + x^3)^2 mod (x^4 + 1, 113) = [94, 109, 11, 18]
    Q, M = 113, 8
    psi = portapi.lib().orc_root_of_unity(M, Q)
    o = portapi.Oracle(k=1, N=4, Q=Q, psi=psi, q=7, qKS=7, n=1, baseG_log2=2, digitsG=3, baseKS=2, dKS=1, variant=1, scheme=1)
    a = np.array([1, 2, 4, 1], dtype=np.uint32)
    A = o.ntt(a, False).astype(np.uint64)
    ab = ((A * A) % Q).astype(np.uint32)
    assert o.ntt(ab, True).tolist() == [94, 109, 11, 18]


def test_round_trip_and_monomial():
    # UnitTestNTT.cpp:53-81 (SwitchFormat twice = identity) + the monomial table definition (mk-cryptoparameters.cpp:53-70)
    g = load_golden("toy_mkntru_k4_N256")
    p = g["p"]
    o = make_oracle(p)
    x = np.random.default_rng(7).integers(0, p.Q, size=(5, p.N), dtype=np.uint32)
    assert (o.ntt(o.ntt(x, False), True) == x).all()
    for m in (0, 1, p.N - 1, p.N, 2 * p.N - 1):
        mono = o.ntt(o.monomial(m), True)
        exp = np.zeros(p.N, dtype=np.int64)
        exp[0] -= 1
        exp[m % p.N] += 1 if m < p.N else -1
        assert (mono == (exp % p.Q).astype(np.uint32)).all()


def test_transpose_is_coefficient_reversal():
    # Transpose + INTT == [a0, -a_{N-1}, ..., -a_1]
    g = load_golden("toy_mklwe_k2_N128")
    p = g["p"]
    o = make_oracle(p)
    a = np.random.default_rng(3).integers(0, p.Q, size=p.N, dtype=np.uint32)
    a[5] = 0
    got = o.transpose_intt(o.ntt(a, False))
    exp = np.concatenate([a[:1], ((p.Q - a[:0:-1].astype(np.int64)) % p.Q).astype(np.uint32)])
    assert (got == exp).all()


needs_ref = pytest.mark.skipif(not refapi.available(), reason="oracle/_ref/libmkfhe_ref.so not built")


@needs_ref
@pytest.mark.parametrize("args", [
    (0, 2, 27, 64, 5, 45181, 45181, 0.75, 32, 1 << 9, 32),
    (1, 2, 27, 64, 5, 32749, 32749, 1.9, 32, 1 << 9, 2),
    (0, 2, 27, 128, 3, 45181, 45181, 0.5, 32, 1 << 6, 32),
])
def test_oracle_matches_compiled_reference_fresh_keys(args):
    r = refapi.Ref(*args)
    r.keygen()
    p = params_from_ref(r.p)
    o = make_oracle(p)
    rng = np.random.default_rng(11)
    bk, pk = r.bk(), r.pkey()
    acc = rng.integers(0, p.Q, size=p.k * p.N, dtype=np.uint32)
    rot = rng.integers(0, p.rot_mod, size=p.k * p.n, dtype=np.uint32)
    assert (r.eval_acc(acc, rot, p.rot_mod) == o.eval_acc(bk, pk, acc, rot, p.rot_mod)).all()
    if p.scheme == 0:
        kA, bad = r.ksk_mntru()
        assert bad == 0  # KSK2[u][j] == j * KSK2[u][1] (mntru-pke.cpp:743-752)
        kB, ctn = None, r.ctnand()
    else:
        (kA, kB), ctn = r.ksk_mklwe(), None
    sk = r.sk()
    for m0, m1 in ((0, 1), (1, 1)):
        c1, b1 = r.encrypt(m0)
        c2, b2 = r.encrypt(m1)
        eo, eb = r.eval_gate(c1, b1, c2, b2)
        oo, ob = o.eval_nand(bk, pk, kA, kB, ctn, c1, b1, c2, b2)
        assert (eo == oo).all() and eb == ob
        assert r.decrypt(eo, eb) == o.decrypt(sk, oo, ob)


@needs_ref
@pytest.mark.slow
def test_oracle_matches_compiled_reference_full_size_eval_acc_prefix():
    """Full-size parameters (STD100_MKNTRU_LWE shape) on random keys of the right shape is covered on the GPU box; here a
    cheap full-N check: NTT / decomposition at N = 2048 against the compiled reference."""
    r = refapi.Ref(1, 2, 27, 4096, 3, 32749, 32749, 1.9, 32, 1 << 9, 2)
    p = params_from_ref(r.p)
    assert (p.Q, p.psi) == (134176769, 100530)
    o = make_oracle(p)
    x = np.random.default_rng(5).integers(0, p.Q, size=(2, p.N), dtype=np.uint32)
    assert (o.ntt(x, False) == r.ntt(x, False)).all()
    assert (o.ntt(x, True) == r.ntt(x, True)).all()
    assert (o.decompose(x[0]) == r.decompose(x[0])).all()


@pytest.mark.skipif(not refapi.available(), reason="oracle/_ref/libmkfhe_ref.so not present")
@pytest.mark.parametrize("scheme", [0, 1])
def test_oracle_keygen_reproduces_reference_keys(scheme):
    """Key generation: the randomness recovered from the reference's keys + secrets, fed to the oracle's
    restatement of KeyGenXZW / KDMKeyGenXZW / Pkey, gives the reference's key bytes back; the recovered messages are the ones
    KeyGenAcc derives from the LWE secret."""
    from keygen_util import expected_messages, extract_randomness, slot_inverse

    r = refapi.Ref(0, 2, 27, 512, 8, 45181, 45181, 0.75, 32, 1 << 9, 32) if scheme == 0 else \
        refapi.Ref(1, 3, 27, 128, 6, 32749, 32749, 1.9, 32, 1 << 7, 2)
    r.keygen()
    p = params_from_ref(r.p)
    o = make_oracle(p)
    bk, pkey, f, crs = r.bk(), r.pkey(), r.unienc_sk(), r.crs()
    x = extract_randomness(p, o, bk, pkey, f, crs)
    assert (x["m"] == expected_messages(p, r.sk())).all()
    inv = slot_inverse(f, p.Q)
    ell = p.digitsG - 1
    ref_bk = bk.reshape(-1, ell, 2, p.N)
    for e in range(ref_bk.shape[0]):
        got = o.keygen_evk(inv[x["party"][e]], crs, x["r"][e], x["e"][e], int(x["m"][e]), bool(x["kdm"][e]))
        assert (got == ref_bk[e]).all(), f"evk {e}"
    ref_pk = pkey.reshape(p.k, ell, p.N)
    for u in range(p.k):
        assert (o.keygen_pkey(f[u], crs, x["pk_e"][u]) == ref_pk[u]).all()
    # the r != 0 branch (the reference's DggR has sigma 0.15: r is 0 in practice): linearity in r of the restated formula
    rng = np.random.default_rng(7)
    rr = rng.integers(0, p.Q, p.N, dtype=np.uint32)
    a = o.keygen_evk(inv[0], crs, rr, x["e"][0], 0, False).astype(np.int64)
    b = o.keygen_evk(inv[0], crs, np.zeros(p.N, np.uint32), x["e"][0], 0, False).astype(np.int64)
    R = o.ntt(rr, False).astype(np.int64)
    g = 1
    for j in range(ell):
        g = g * (1 << p.baseG_log2) % p.Q
        assert ((a[j, 1] - b[j, 1]) % p.Q == g * R % p.Q * inv[0].astype(np.int64) % p.Q).all()
        assert ((a[j, 0] - b[j, 0]) % p.Q == R[j] * crs[j].astype(np.int64) % p.Q).all()
    r.close()


@pytest.mark.skipif(not refapi.available(), reason="oracle/_ref/libmkfhe_ref.so not present")
def test_oracle_ksk_keygen_reproduces_reference_keys():
    """Key-switch key generation: noise recovered from the reference's keys and secrets, fed to the
    oracle's restatement of KeySwitchGen2 (digit-1 slice) / MKLWE KeySwitchGen, gives the reference's key bytes back."""
    from keygen_util import extract_ksk_mklwe_noise, extract_ksk_mntru_noise, switch_modulus

    r = refapi.Ref(0, 2, 27, 512, 8, 45181, 45181, 0.75, 32, 1 << 9, 32)
    r.keygen()
    p = params_from_ref(r.p)
    o = make_oracle(p)
    ksk, bad = r.ksk_mntru()
    assert bad == 0
    F, Fi = r.mntru_F(False), r.mntru_F(True)
    s_q = switch_modulus(r.unienc_sk_coef(), p.Q, p.qKS)
    E = extract_ksk_mntru_noise(p, ksk, F, s_q)
    ksk = ksk.reshape(p.k, p.N * p.dKS, p.n)
    for u in range(p.k):
        assert (o.keygen_ksk_mntru(s_q[u], E[u], Fi[u]) == ksk[u]).all()
    r.close()

    r = refapi.Ref(1, 3, 27, 128, 6, 32749, 32749, 1.9, 32, 1 << 7, 2)
    r.keygen()
    p = params_from_ref(r.p)
    o = make_oracle(p)
    A, B = r.ksk_mklwe()
    s = r.sk().reshape(p.k, p.n)
    sN_q = switch_modulus(r.unienc_sk_coef(), p.Q, p.qKS)
    e = extract_ksk_mklwe_noise(p, A, B, s, sN_q)
    rows = p.N * p.baseKS * p.dKS
    A, B = A.reshape(p.k, rows, p.n), B.reshape(p.k, rows)
    for u in range(p.k):
        assert (o.keygen_ksk_mklwe(A[u], e[u], s[u], sN_q[u]) == B[u]).all()
    r.close()
