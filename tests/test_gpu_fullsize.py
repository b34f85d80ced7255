"""GPU: full-size parameter sets (the configs of BASELINE.json) against the compiled reference (oracle/_ref, prebuilt; it
generates the keys on the host exactly as the reference's examples do) and the C oracle; plus the drop-in test: the
reference's own EvalBinGate orchestration running on the GPU accumulator through integration/unienc_acc_b200.h.
"""
import numpy as np
import pytest

from oracle import refapi
from util import make_oracle, params_from_ref

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not refapi.available(), reason="oracle/_ref/libmkfhe_ref.so not present")]


@pytest.fixture(scope="module", params=["STD100_MKNTRU_LWE", "STD100_MKNTRU"])
def full(request):
    from gpu_util import ctx_from_params, load_ref_keys

    r = refapi.Ref.from_paramset(request.param)
    r.keygen()
    p = params_from_ref(r.p)
    ctx = ctx_from_params(p)
    keys = load_ref_keys(ctx, r)
    yield r, p, ctx, keys
    ctx.close()
    r.close()


def _encrypt_batch(r, bits0, bits1):
    c1, b1, c2, b2 = [], [], [], []
    for m0, m1 in zip(bits0, bits1):
        x, bx = r.encrypt(int(m0))
        y, by = r.encrypt(int(m1))
        c1.append(x); b1.append(bx); c2.append(y); b2.append(by)
    P = r.p
    return (np.array(c1, dtype=np.uint32).reshape(-1, P.k, P.n), np.array(b1, dtype=np.uint32),
            np.array(c2, dtype=np.uint32).reshape(-1, P.k, P.n), np.array(b2, dtype=np.uint32))


def test_gate_batch_vs_reference_and_nand_truth(full):
    r, p, ctx, keys = full
    rng = np.random.default_rng(2)
    B = 48
    m0, m1 = rng.integers(0, 2, B), rng.integers(0, 2, B)
    c1, b1, c2, b2 = _encrypt_batch(r, m0, m1)
    if p.scheme == 0:
        out = ctx.EvalBinGate("NAND", c1, c2)
        bout = np.zeros(B, dtype=np.uint32)
    else:
        out, bout = ctx.EvalBinGate("NAND", c1, c2, b1, b2)
    # identical refreshed ciphertexts on a sample (the reference costs ~1 s per gate on the CPU)
    for i in range(3):
        eo, eb = r.eval_gate(c1[i].ravel(), b1[i], c2[i].ravel(), b2[i])
        assert (out[i].ravel() == eo).all() and int(bout[i]) == eb
    # every gate decrypts to NAND (size-independent property)
    o = make_oracle(p)
    bits = np.array([o.decrypt(keys["sk"], out[i].ravel(), int(bout[i])) for i in range(B)])
    assert (bits == 1 - (m0 & m1)).all()
    assert all(r.decrypt(out[i].ravel(), int(bout[i])) == bits[i] for i in range(4))
    # the GPU's batched Decrypt agrees with the oracle on every gate and with the reference's Decrypt on the sample
    ctx.LoadSecretKey(keys["sk"])
    assert (ctx.Decrypt(out, bout if p.scheme == 1 else None) == bits).all()


def test_eval_acc_vs_reference(full):
    r, p, ctx, keys = full
    rng = np.random.default_rng(3)
    acc = rng.integers(0, p.Q, size=(2, p.k, p.N), dtype=np.uint32)
    rot = rng.integers(0, p.rot_mod, size=(2, p.k, p.n), dtype=np.uint32)
    got = ctx.EvalAcc(acc, rot, p.rot_mod)
    assert (got[0].ravel() == r.eval_acc(acc[0].ravel(), rot[0].ravel(), p.rot_mod)).all()
    o = make_oracle(p)
    assert (got[1].ravel() == o.eval_acc(keys["bk"], keys["pkey"], acc[1].ravel(), rot[1].ravel(), p.rot_mod)).all()


def test_ntt_full_size(full):
    r, p, ctx, keys = full
    x = np.random.default_rng(5).integers(0, p.Q, size=(64, p.N), dtype=np.uint32)
    f = ctx.NTT(x)
    assert (f[:8] == r.ntt(x[:8], False)).all()
    assert (ctx.NTT(f, inverse=True) == x).all()


@pytest.mark.parametrize("B", [74, 148, 200, 400, 1200])
def test_identical_gates_give_identical_outputs(full, B):
    """Race / scheduling check (compute-sanitizer is closed on the GPU pool): 1200 copies of one refreshed gate -- more than two
    waves of resident CTAs, so copies run at different times, on different SMs and next to different neighbours -- must give
    1200 identical refreshed ciphertexts, equal to the reference's.  74 copies: every SM pair of the GPU runs one gate as a
    thread-block cluster at once (remote shared-memory writes and cluster barriers of 74 clusters in flight).  148 copies: the same on the party-parallel kernel that
    serves batches up to the SM count (named barriers between groups, shared-memory atomics for sumV); 200 and 400 copies: the
    128- and 80-register builds of the one-group kernel, which the host picks when fewer gates than the budget share an SM."""
    r, p, ctx, keys = full
    c1, b1 = r.encrypt(1)
    c2, b2 = r.encrypt(0)
    c3, b3 = r.encrypt(1)
    c4, b4 = r.encrypt(1)
    x, bx = r.eval_gate(c1, b1, c2, b2)   # refreshed inputs: dense rotation amounts
    y, by = r.eval_gate(c3, b3, c4, b4)
    eo, eb = r.eval_gate(x, bx, y, by)
    X = np.tile(np.asarray(x, dtype=np.uint32).reshape(1, p.k, p.n), (B, 1, 1))
    Y = np.tile(np.asarray(y, dtype=np.uint32).reshape(1, p.k, p.n), (B, 1, 1))
    if p.scheme == 1:
        out, bout = ctx.EvalBinGate("NAND", X, Y, np.full(B, bx, np.uint32), np.full(B, by, np.uint32))
        assert (bout == eb).all()
    else:
        out = ctx.EvalBinGate("NAND", X, Y)
    out = np.asarray(out).reshape(B, -1)
    assert (out == out[0]).all()
    assert (out[0] == eo).all()


def test_ntt_16_coefficient_variant(full, monkeypatch):
    """The experimental two-exchange transform (16 coefficients per thread, MKFHE_NTT16=1) against the reference."""
    r, p, ctx, keys = full
    monkeypatch.setenv("MKFHE_NTT16", "1")
    x = np.random.default_rng(6).integers(0, p.Q, size=(33, p.N), dtype=np.uint32)
    f = ctx.NTT(x)
    assert (f[:4] == r.ntt(x[:4], False)).all()
    assert (ctx.NTT(x, inverse=True)[:4] == r.ntt(x[:4], True)).all()
    assert (ctx.NTT(f, inverse=True) == x).all()


def test_drop_in_accumulator_through_reference_api(full):
    """The reference's unmodified BinFHEScheme::EvalBinGate with UniEncAccumulatorB200 plugged into UniEncACCscheme."""
    from mkfhe_b200.lib import LIB_PATH

    r, p, ctx, keys = full
    c1, b1 = r.encrypt(1)
    c2, b2 = r.encrypt(1)
    cpu_out, cpu_b = r.eval_gate(c1, b1, c2, b2)
    r.use_accumulator_b200(LIB_PATH)
    try:
        gpu_out, gpu_b = r.eval_gate(c1, b1, c2, b2)
    finally:
        r.use_accumulator_b200(LIB_PATH, restore=True)
    assert (cpu_out == gpu_out).all() and cpu_b == gpu_b
    assert r.decrypt(gpu_out, gpu_b) == 0


def test_adder_circuit_decrypts_to_the_sum(full):
    """SURVEY 8f rank 4, "gate circuits on the batch API": 16 independent 4-bit ripple-carry adders (NAND only, 576 gates)
    through mkfhe_eval_circuit -- one batch per level (12 launches of the gate kernel) instead of the reference's one
    EvalBinGate per gate.  Every sum, decrypted by the reference's Decrypt, equals a + b (MK-LWE flavour); the first XOR
    (4 gates, 3 levels) equals the reference's own EvalBinGate chain bit for bit (both flavours)."""
    r, p, ctx, keys = full
    rng = np.random.default_rng(17)
    bits, inst = 4, 16
    avals, bvals = rng.integers(0, 16, inst), rng.integers(0, 16, inst)
    from mkfhe_b200.netlist import ripple_carry_adders

    per = 2 * bits + 1  # a bits | b bits | carry-in (an encryption of 0)
    nl, sums = ripple_carry_adders(bits, inst)
    gates, w = nl.gates, nl.n_wires
    assert len(gates) == 9 * bits * inst
    wires = np.zeros((w, p.k, p.n), dtype=np.uint32)
    wb = np.zeros(w, dtype=np.uint32)
    for t in range(inst):
        plain = [(int(avals[t]) >> i) & 1 for i in range(bits)] + [(int(bvals[t]) >> i) & 1 for i in range(bits)] + [0]
        for i, m in enumerate(plain):
            ct, b = r.encrypt(m)
            wires[per * t + i] = np.asarray(ct, dtype=np.uint32).reshape(p.k, p.n)
            wb[per * t + i] = b
    out, ob, levels = ctx.EvalCircuit(wires, np.array(gates, dtype=np.uint32), wb if p.scheme == 1 else None)
    if ob is None:
        ob = np.zeros(w, dtype=np.uint32)
    assert levels == 2 * bits + 4  # the carry chain
    # The reference's MKNTRU gate outputs do not compose: in the reference itself NAND(1, NAND(0, 1)) decrypts to 1 (checked on
    # the CPU with its own Encrypt / EvalBinGate / Decrypt), so only the MK-LWE flavour has a meaningful sum.  Parity with the
    # reference's chain (below) holds for both.
    for t in range(inst if p.scheme == 1 else 0):
        got = sum(r.decrypt(out[s].ravel(), int(ob[s])) << i for i, s in enumerate(sums[t]))
        assert got == int(avals[t]) + int(bvals[t])
    # the first adder's bit-0 XOR against the reference, gate by gate
    ew = {0: (wires[0].ravel(), int(wb[0])), bits: (wires[bits].ravel(), int(wb[bits]))}
    for i1, i2, o in gates[:4]:
        ew[o] = r.eval_gate(ew[i1][0], ew[i1][1], ew[i2][0], ew[i2][1])
        assert (out[o].ravel() == ew[o][0]).all() and int(ob[o]) == int(ew[o][1])


def test_batched_gate_class_over_reference_objects(full):
    """integration/binfhe_batch_b200.h: the C++ EvalBinGateBatch a maintainer adds, compiled against the unmodified reference.
    It takes the reference's BinFHECryptoParams + UniEncBTKey (+ ctNAND), flattens and uploads them itself, takes vectors of
    the reference's ciphertext objects and returns ciphertext objects: every output equals the reference's own EvalBinGate on
    the same pair, the reference's Decrypt gives NAND, and the argument checks throw the reference's config_error."""
    from mkfhe_b200.lib import LIB_PATH

    r, p, ctx, keys = full
    bits = [(0, 0), (0, 1), (1, 0), (1, 1), (1, 1), (0, 1)]
    enc = [(r.encrypt(x), r.encrypt(y)) for x, y in bits]
    ct1 = np.stack([e[0][0] for e in enc]).reshape(len(bits), p.k, p.n)
    ct2 = np.stack([e[1][0] for e in enc]).reshape(len(bits), p.k, p.n)
    b1 = np.array([e[0][1] for e in enc], dtype=np.uint32)
    b2 = np.array([e[1][1] for e in enc], dtype=np.uint32)
    r.batch_b200_load(LIB_PATH)
    out, bout, dec = r.batch_b200_gates(ct1, b1, ct2, b2)
    assert list(dec) == [1 - (x & y) for x, y in bits]
    for g in (0, 3, 5):  # the reference takes 1-2 s a gate
        eo, eb = r.eval_gate(ct1[g].ravel(), int(b1[g]), ct2[g].ravel(), int(b2[g]))
        assert (out[g].ravel() == eo).all() and int(bout[g]) == int(eb)
    assert r.batch_b200_errors() == 15  # wrong gate | one object twice | ragged vectors | a circuit input that holds no ciphertext
    # EvalCircuit of the same class: a + b for one 2-bit adder on reference ciphertext objects
    from mkfhe_b200.netlist import ripple_carry_adders

    nl, sums = ripple_carry_adders(2, 1)
    plain = [1, 1, 1, 0, 0]  # a = 3, b = 1, carry-in 0
    wires = np.zeros((nl.n_wires, p.k, p.n), dtype=np.uint32)
    wb = np.zeros(nl.n_wires, dtype=np.uint32)
    for i, m in enumerate(plain):
        ct, b = r.encrypt(m)
        wires[i], wb[i] = np.asarray(ct, dtype=np.uint32).reshape(p.k, p.n), b
    cw, cb, levels, cdec = r.batch_b200_circuit(wires, wb, len(plain), nl.array())
    gw, gb, glev = ctx.EvalCircuit(wires, nl.array(), wb if p.scheme == 1 else None)
    assert levels == glev == 8 and (cw == gw).all() and (p.scheme == 0 or (cb == gb).all())
    if p.scheme == 1:  # MKNTRU outputs do not compose in the reference (see test_adder_circuit_decrypts_to_the_sum)
        assert sum(int(cdec[s]) << i for i, s in enumerate(sums[0])) == 4


@pytest.mark.gpu
@pytest.mark.parametrize("batch", [1, 2, 7, 64])
def test_ntt_n1024_vs_oracle(batch):
    """Config 5 at N = 1024 (Q = 134215681, psi = 282116: SURVEY.md 8c): the batched transform runs two polynomials per
    2048-word pass of the specialised kernel (an odd last one on the generic kernel); bit-exact against the C oracle, whose
    transform is pinned to NativePoly::SetFormat (tests/test_oracle.py)."""
    from types import SimpleNamespace

    from gpu_util import ctx_from_params
    from util import make_oracle

    p = SimpleNamespace(k=2, N=1024, Q=134215681, psi=282116, q=32749, qKS=32749, n=16, baseG_log2=9, digitsG=3, baseKS=32, dKS=3,
                        variant=1, scheme=1)
    ctx = ctx_from_params(p)
    o = make_oracle(p)
    x = np.random.default_rng(50 + batch).integers(0, p.Q, size=(batch, p.N), dtype=np.uint32)
    x[0, :4] = [0, 1, p.Q - 1, p.Q - 2]
    f = ctx.NTT(x)
    assert (f == o.ntt(x, False)).all()
    assert (ctx.NTT(x, inverse=True) == o.ntt(x, True)).all()
    assert (ctx.NTT(f, inverse=True) == x).all()
    ctx.close()
