"""GPU: bootstrapping-key and Pkey generation against the reference's own keys.

The randomness the reference sampled is recovered from its keys and secrets (tests/keygen_util.py); the GPU generator fed
with it must return the reference's key BYTES, at toy size (both schemes, odd gadget base) and at the full boolean-mklwe
size, and the keys it installs with load=True must bootstrap to the reference's EvalBinGate output.  The r != 0 branch (the
reference's DggR has sigma 0.15, so it samples r = 0 in practice) is checked against the C oracle's restatement.
"""
import numpy as np
import pytest

from oracle import refapi
from util import make_oracle, params_from_ref

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not refapi.available(), reason="oracle/_ref/libmkfhe_ref.so not present")]

TOY = {
    "mkntru_k2_N256": (0, 2, 27, 512, 8, 45181, 45181, 0.75, 32, 1 << 9, 32),
    "mklwe_k3_N64_B7": (1, 3, 27, 128, 6, 32749, 32749, 1.9, 32, 1 << 7, 2),
}


def _setup(r):
    from gpu_util import ctx_from_params
    from keygen_util import extract_randomness, slot_inverse

    r.keygen()
    p = params_from_ref(r.p)
    o = make_oracle(p)
    bk, pkey, f, crs = r.bk(), r.pkey(), r.unienc_sk(), r.crs()
    x = extract_randomness(p, o, bk, pkey, f, crs)
    return p, o, ctx_from_params(p), bk, pkey, f, crs, slot_inverse(f, p.Q), x


@pytest.mark.parametrize("name", sorted(TOY))
def test_keygen_reproduces_reference_keys_toy(name):
    r = refapi.Ref(*TOY[name])
    p, o, ctx, bk, pkey, f, crs, inv, x = _setup(r)
    assert (ctx.KeyGenAcc(inv, crs, x["r"], x["e"], x["m"]) == bk).all()
    assert (ctx.KeyGenPkey(f, crs, x["pk_e"]) == pkey).all()
    # r != 0: against the oracle's restatement (pinned to the reference for r = 0 and by linearity in tests/test_oracle.py)
    rng = np.random.default_rng(11)
    rr = x["r"].copy()
    rr[:] = rng.integers(0, p.Q, size=rr.shape, dtype=np.uint32)
    got = ctx.KeyGenAcc(inv, crs, rr, x["e"], x["m"]).reshape(x["e"].shape)
    for e in (0, 1, got.shape[0] // 2, got.shape[0] - 1):
        assert (got[e] == o.keygen_evk(inv[x["party"][e]], crs, rr[e], x["e"][e], int(x["m"][e]), bool(x["kdm"][e]))).all()
    ctx.close()
    r.close()


def test_keygen_full_size_and_bootstraps():
    """boolean-mklwe (STD100_MKNTRU_LWE): 1001 evks of 32 KiB regenerated on the GPU = the reference's 31 MiB key; installed
    with load=True (plus the reference's key-switch key) the context bootstraps to the reference's EvalBinGate output."""
    r = refapi.Ref.from_paramset("STD100_MKNTRU_LWE")
    p, o, ctx, bk, pkey, f, crs, inv, x = _setup(r)
    assert (ctx.KeyGenAcc(inv, crs, x["r"], x["e"], x["m"], load=True) == bk).all()
    assert (ctx.KeyGenPkey(f, crs, x["pk_e"], load=True) == pkey).all()
    from mkfhe_b200.context import _ptr

    kskA, kskB = r.ksk_mklwe()
    (pa, na), (pb, nb) = _ptr(kskA), _ptr(kskB)
    ctx._ck(ctx._L.mkfhe_load_ksk_mklwe(ctx._h, pa, na, pb, nb), "mkfhe_load_ksk_mklwe")

    def both(x1, y1, x2, y2):  # the same gate on the reference and on the GPU
        eo, eb = r.eval_gate(x1, y1, x2, y2)
        out, bout = ctx.EvalBinGate("NAND", x1.reshape(1, p.k, p.n), x2.reshape(1, p.k, p.n), np.array([y1], np.uint32),
                                    np.array([y2], np.uint32))
        assert (out.ravel() == eo).all() and int(bout[0]) == eb
        return eo, eb

    (c1, b1), (c2, b2), (c3, b3), (c4, b4) = r.encrypt(1), r.encrypt(0), r.encrypt(1), r.encrypt(1)
    o1, ob1 = both(c1, b1, c2, b2)
    o2, ob2 = both(c3, b3, c4, b4)
    o3, ob3 = both(o1, ob1, o2, ob2)  # refreshed inputs: dense rotations, every step of the blind rotation runs
    assert r.decrypt(o3, ob3) == 1    # NAND(NAND(1,0), NAND(1,1)) = NAND(1, 0) = 1
    ctx.close()
    r.close()


def test_ksk_keygen_reproduces_reference_keys():
    """Key-switch key generation at toy size, both schemes: the GPU fed with the noise recovered from the
    reference's keys returns the reference's key bytes; with perturbed (non-replicated) noise it matches the C oracle."""
    from gpu_util import ctx_from_params
    from keygen_util import extract_ksk_mklwe_noise, extract_ksk_mntru_noise, switch_modulus

    r = refapi.Ref(*TOY["mkntru_k2_N256"])
    r.keygen()
    p = params_from_ref(r.p)
    o, ctx = make_oracle(p), ctx_from_params(p)
    ksk, bad = r.ksk_mntru()
    F, Fi = r.mntru_F(False), r.mntru_F(True)
    s_q = switch_modulus(r.unienc_sk_coef(), p.Q, p.qKS)
    E = extract_ksk_mntru_noise(p, ksk, F, s_q)
    assert (ctx.KeySwitchGen2(s_q, E, Fi) == ksk).all()
    E2 = np.random.default_rng(3).integers(0, p.qKS, size=E.shape, dtype=np.uint32)
    got = ctx.KeySwitchGen2(s_q, E2, Fi).reshape(p.k, -1, p.n)
    assert (got[1] == o.keygen_ksk_mntru(s_q[1], E2[1], Fi[1])).all()
    ctx.close()
    r.close()

    r = refapi.Ref(*TOY["mklwe_k3_N64_B7"])
    r.keygen()
    p = params_from_ref(r.p)
    o, ctx = make_oracle(p), ctx_from_params(p)
    A, B = r.ksk_mklwe()
    s = r.sk().reshape(p.k, p.n)
    sN_q = switch_modulus(r.unienc_sk_coef(), p.Q, p.qKS)
    e = extract_ksk_mklwe_noise(p, A, B, s, sN_q)
    assert (ctx.KeySwitchGenLWE(A, e, s, sN_q) == B).all()
    A2 = np.random.default_rng(4).integers(0, p.qKS, size=A.shape, dtype=np.uint32)
    rows = p.N * p.baseKS * p.dKS
    got = ctx.KeySwitchGenLWE(A2, e, s, sN_q).reshape(p.k, rows)
    assert (got[2] == o.keygen_ksk_mklwe(A2.reshape(p.k, rows, p.n)[2], e[2], s[2], sN_q[2])).all()
    ctx.close()
    r.close()


def test_ksk_keygen_full_size_mntru():
    """boolean-mkntru (STD100_MKNTRU): the 2 x 8192 x 560 digit-1 slice regenerated on the GPU = the reference's; installed with
    load=True next to the reference's other keys it bootstraps to the reference's EvalBinGate output."""
    from gpu_util import ctx_from_params
    from keygen_util import extract_ksk_mntru_noise, switch_modulus

    r = refapi.Ref.from_paramset("STD100_MKNTRU")
    r.keygen()
    p = params_from_ref(r.p)
    ctx = ctx_from_params(p)
    ksk, bad = r.ksk_mntru()
    assert bad == 0
    F, Fi = r.mntru_F(False), r.mntru_F(True)
    s_q = switch_modulus(r.unienc_sk_coef(), p.Q, p.qKS)
    E = extract_ksk_mntru_noise(p, ksk, F, s_q)
    ctx.LoadKeys(r.bk(), r.pkey(), ksk=np.zeros_like(ksk), ctnand=r.ctnand())  # a wrong key-switch key first
    assert (ctx.KeySwitchGen2(s_q, E, Fi, load=True) == ksk).all()
    c1, _ = r.encrypt(1)
    c2, _ = r.encrypt(1)
    eo, _ = r.eval_gate(c1, 0, c2, 0)
    out = ctx.EvalBinGate("NAND", c1.reshape(1, p.k, p.n), c2.reshape(1, p.k, p.n))
    assert (np.asarray(out).ravel() == eo).all() and r.decrypt(eo, 0) == 0
    ctx.close()
    r.close()


@pytest.mark.parametrize("name", sorted(TOY))
def test_encrypt_batch_reproduces_reference_ciphertexts(name):
    """Batched Encrypt: the noise of ciphertexts the reference encrypted is recovered with the secret
    (MNTRU: e = c x F - message; MKLWE: err = b - message - <a, s>), and the GPU fed with it returns the reference's
    ciphertexts; random noise is checked against the numpy restatement."""
    from gpu_util import ctx_from_params
    from keygen_util import _cen, encrypt_mklwe_np, encrypt_mntru_np

    r = refapi.Ref(*TOY[name])
    r.keygen()
    p = params_from_ref(r.p)
    ctx = ctx_from_params(p)
    q = p.q
    bits = np.array([0, 1, 1, 0, 1, 0, 0, 1, 1], dtype=np.uint8)
    cts = [r.encrypt(int(b)) for b in bits]
    ct = np.array([c for c, _ in cts], dtype=np.uint32).reshape(len(bits), p.k, p.n)
    bs = np.array([b for _, b in cts], dtype=np.uint32)
    rng = np.random.default_rng(12)
    if p.scheme == 0:
        F, Fi = r.mntru_F(False).astype(np.int64), r.mntru_F(True)
        e = np.stack([(ct[:, u, :].astype(np.int64) @ F[u]) % q for u in range(p.k)], axis=1)
        e[:, 0, 0] = (e[:, 0, 0] - bits.astype(np.int64) * (q // 4)) % q
        assert np.abs(_cen(e, q)).max() < 16
        assert (ctx.EncryptBatch(bits, e, Fi) == ct).all()
        e2 = rng.integers(0, q, size=(70, p.k, p.n), dtype=np.uint32)
        m2 = rng.integers(0, 2, size=70, dtype=np.uint8)
        assert (ctx.EncryptBatch(m2, e2, Fi) == encrypt_mntru_np(p, Fi, e2, m2)).all()
    else:
        s = r.sk().reshape(p.k, p.n)
        err = (bs.astype(np.int64) - bits.astype(np.int64) * (q // 4)
               - ct.reshape(len(bits), -1).astype(np.int64) @ s.reshape(-1).astype(np.int64)) % q
        assert np.abs(_cen(err, q)).max() < 16
        assert (ctx.EncryptBatch(bits, err, s, a=ct) == bs).all()
        a2 = rng.integers(0, q, size=(70, p.k, p.n), dtype=np.uint32)
        err2 = rng.integers(0, q, size=70, dtype=np.uint32)
        m2 = rng.integers(0, 2, size=70, dtype=np.uint8)
        assert (ctx.EncryptBatch(m2, err2, s, a=a2) == encrypt_mklwe_np(p, s, a2, err2, m2)).all()
    ctx.close()
    r.close()
