"""GPU: the CUDA path, called through the C ABI, against the golden fixtures (compiled-reference outputs) and the oracle.
Bit-exact: every comparison is == on uint32 arrays.
"""
import numpy as np
import pytest

from util import golden_names, load_golden, make_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=golden_names())
def fx(request):
    from gpu_util import ctx_from_params, load_fixture_keys

    g = load_golden(request.param)
    ctx = ctx_from_params(g["p"])
    load_fixture_keys(ctx, g)
    yield g, ctx
    ctx.close()


def test_ntt_matches_reference(fx):
    g, ctx = fx
    assert (ctx.NTT(g["ntt_in"], inverse=False) == g["ntt_fwd"]).all()
    assert (ctx.NTT(g["ntt_in"], inverse=True) == g["ntt_inv"]).all()
    assert (ctx.NTT(ctx.NTT(g["ntt_in"]), inverse=True) == g["ntt_in"]).all()


def test_decompose_matches_reference(fx):
    g, ctx = fx
    assert (ctx.SignedDigitDecompose(g["dec_in"]) == g["dec_out"]).all()


def test_eval_acc_matches_reference(fx):
    g, ctx = fx
    p = g["p"]
    got = ctx.EvalAcc(g["acc_in"].reshape(-1, p.k, p.N), g["acc_rot"].reshape(-1, p.k, p.n), p.rot_mod)
    assert (got.reshape(g["acc_out"].shape) == g["acc_out"]).all()


def test_eval_bin_gate_matches_reference(fx):
    g, ctx = fx
    p = g["p"]
    if p.scheme == 0:
        out = ctx.EvalBinGate("NAND", g["g_ct1"], g["g_ct2"])
        assert (out == g["g_out"]).all()
    else:
        out, bout = ctx.EvalBinGate("NAND", g["g_ct1"], g["g_ct2"], g["g_b1"], g["g_b2"])
        assert (out == g["g_out"]).all() and (bout == g["g_bout"]).all()
    o = make_oracle(p)
    bits = [o.decrypt(g["sk"], out[i], int(g["g_bout"][i])) for i in range(len(out))]
    assert bits == [int(b) for b in g["g_bits"]]


def test_decrypt_matches_reference(fx):
    """batched Decrypt: the reference's plaintexts of its own gate outputs (g_bits), the oracle on random ciphertexts
    (every residue class of the rounding), and the error paths"""
    import mkfhe_b200 as mk

    g, ctx = fx
    p = g["p"]
    lwe = p.scheme == 1
    with pytest.raises(mk.MkfheError):  # no secret key yet
        ctx.Decrypt(g["g_out"].reshape(-1, p.k, p.n), g["g_bout"] if lwe else None)
    ctx.LoadSecretKey(g["sk"])
    got = ctx.Decrypt(g["g_out"].reshape(-1, p.k, p.n), g["g_bout"] if lwe else None)
    assert (got == g["g_bits"]).all()
    o = make_oracle(p)
    rng = np.random.default_rng(9)
    B = 257  # ragged against the 8-ciphertexts-per-block launch
    ct = rng.integers(0, p.q, size=(B, p.k, p.n), dtype=np.uint32)
    b = rng.integers(0, p.q, size=B, dtype=np.uint32)
    want = np.array([o.decrypt(g["sk"], ct[i].ravel(), int(b[i])) for i in range(B)], dtype=np.uint32)
    assert (ctx.Decrypt(ct, b if lwe else None) == want).all()
    assert set(np.unique(want)) == {0, 1, 2, 3}
    assert len(ctx.Decrypt(ct[:0], b[:0] if lwe else None)) == 0
    with pytest.raises(mk.MkfheError):
        ctx.Decrypt(ct[:1], b[:1] if lwe else None, p=1)


def test_eval_acc_random_vs_oracle(fx):
    """fresh random accumulators / rotations (incl. ragged batch sizes) against the C oracle"""
    g, ctx = fx
    p = g["p"]
    o = make_oracle(p)
    rng = np.random.default_rng(42)
    for B in (1, 5):
        acc = rng.integers(0, p.Q, size=(B, p.k, p.N), dtype=np.uint32)
        rot = rng.integers(0, p.rot_mod, size=(B, p.k, p.n), dtype=np.uint32)
        got = ctx.EvalAcc(acc, rot, p.rot_mod)
        for b in range(B):
            exp = o.eval_acc(g["bk"], g["pkey"], acc[b].ravel(), rot[b].ravel(), p.rot_mod)
            assert (got[b].ravel() == exp).all()


def test_errors_and_edge_cases(fx):
    import mkfhe_b200 as mk

    g, ctx = fx
    p = g["p"]
    with pytest.raises(mk.config_error):
        ctx.EvalBinGate("AND", g["g_ct1"], g["g_ct2"], g.get("g_b1"), g.get("g_b2"))
    with pytest.raises(mk.config_error):
        ctx.EvalBinGate("NAND", g["g_ct1"], g["g_ct1"], g.get("g_b1"), g.get("g_b1"))
    # empty batch is a no-op
    e = np.zeros((0, p.k, p.N), dtype=np.uint32)
    assert ctx.EvalAcc(e, np.zeros((0, p.k, p.n), dtype=np.uint32), p.rot_mod).shape == e.shape
    # wrong key size is rejected, not read out of bounds
    with pytest.raises(mk.MkfheError):
        ctx.LoadKeys(g["bk"][:-1], g["pkey"])


@pytest.mark.parametrize("name", ["toy_mkntru_k4_N256", "toy_mklwe_k2_N128"])
def test_unfused_key_switch_path(name, monkeypatch):
    """the stand-alone key-switch kernels (used when qKS >= 2^16 rules out the u16 rows of the fused epilogue)"""
    from gpu_util import ctx_from_params, load_fixture_keys

    monkeypatch.setenv("MKFHE_UNFUSED_KEYSWITCH", "1")
    g = load_golden(name)
    p = g["p"]
    ctx = ctx_from_params(p)
    try:
        load_fixture_keys(ctx, g)
        if p.scheme == 0:
            assert (ctx.EvalBinGate("NAND", g["g_ct1"], g["g_ct2"]) == g["g_out"]).all()
        else:
            out, bout = ctx.EvalBinGate("NAND", g["g_ct1"], g["g_ct2"], g["g_b1"], g["g_b2"])
            assert (out == g["g_out"]).all() and (bout == g["g_bout"]).all()
    finally:
        ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["toy_mkntru_k4_N256", "toy_mklwe_k2_N128"])
def test_keyfile_load_gives_the_same_gates(name, tmp_path):
    """Keys written to the binary key file and loaded with mkfhe_keyfile_load bootstrap exactly like keys loaded from
    memory (= the reference's golden outputs); a file for another parameter set is refused."""
    import mkfhe_b200 as mk
    from gpu_util import ctx_from_params

    g = load_golden(name)
    p = g["p"]
    ctx = ctx_from_params(p)
    path = str(tmp_path / "keys.mkfhe")
    if p.scheme == 0:
        mk.SaveKeyFile(path, ctx.params, g["bk"], g["pkey"], ksk=g["ksk"], ctnand=g["ctnand"])
    else:
        mk.SaveKeyFile(path, ctx.params, g["bk"], g["pkey"], kskA=g["kskA"], kskB=g["kskB"])
    ctx.LoadKeyFile(path)
    if p.scheme == 0:
        out = ctx.EvalBinGate("NAND", g["g_ct1"], g["g_ct2"])
        assert (out == g["g_out"]).all()
    else:
        out, bout = ctx.EvalBinGate("NAND", g["g_ct1"], g["g_ct2"], g["g_b1"], g["g_b2"])
        assert (out == g["g_out"]).all() and (bout == g["g_bout"]).all()
    ctx.close()
    other = load_golden("toy_mklwe_k2_N128" if p.scheme == 0 else "toy_mkntru_k4_N256")
    ctx2 = ctx_from_params(other["p"])
    with pytest.raises(mk.MkfheError):
        ctx2.LoadKeyFile(path)
    ctx2.close()


def _oracle_gate(o, g, x, bx, y, by):
    return o.eval_nand(g["bk"], g["pkey"], g["kskA_any"], g["kskB_any"], g["ctnand_any"], x, bx, y, by)


def test_circuit_matches_gate_by_gate_oracle(fx):
    """mkfhe_eval_circuit: a NAND netlist (XOR of the first two gate inputs, then gates mixing every level) levelised and run
    level by level on the device equals the same netlist evaluated gate by gate by the C oracle, bit for bit on every wire;
    primary inputs stay untouched."""
    g, ctx = fx
    p = g["p"]
    c1 = np.asarray(g["g_ct1"], dtype=np.uint32).reshape(-1, p.k, p.n)
    c2 = np.asarray(g["g_ct2"], dtype=np.uint32).reshape(-1, p.k, p.n)
    lwe = p.scheme == 1
    b1 = np.asarray(g["g_b1"], dtype=np.uint32) if lwe else np.zeros(len(c1), np.uint32)
    b2 = np.asarray(g["g_b2"], dtype=np.uint32) if lwe else np.zeros(len(c1), np.uint32)
    # wires 0, 1: inputs; 2..5: XOR; 6: NAND(xor, w2) (levels 3); 7: NAND(w0, w6); 8: NAND(w3, w4) again (same as 5)
    gates = np.array([(0, 1, 2), (0, 2, 3), (1, 2, 4), (3, 4, 5), (5, 2, 6), (0, 6, 7), (4, 3, 8)], dtype=np.uint32)
    W = 9
    wires = np.zeros((W, p.k, p.n), dtype=np.uint32)
    wb = np.zeros(W, dtype=np.uint32)
    wires[0], wires[1], wb[0], wb[1] = c1[0], c2[0], b1[0], b2[0]
    out, ob, levels = ctx.EvalCircuit(wires, gates, wb if lwe else None)
    assert levels == 5
    assert (out[:2] == wires[:2]).all()
    o = make_oracle(p)
    ew, eb = wires.copy(), wb.copy()
    for i1, i2, oo in gates:
        r = _oracle_gate(o, g, ew[i1].ravel(), int(eb[i1]), ew[i2].ravel(), int(eb[i2]))
        ew[oo], eb[oo] = np.asarray(r[0]).reshape(p.k, p.n), int(r[1])
    assert (out == ew).all()
    if lwe:
        assert (ob == eb).all()
    assert (out[8] == out[5]).all()  # commutativity of the input sum: NAND(w4, w3) == NAND(w3, w4)


def test_circuit_argument_checks(fx):
    import mkfhe_b200 as mk

    g, ctx = fx
    p = g["p"]
    wires = np.zeros((4, p.k, p.n), dtype=np.uint32)
    wb = np.zeros(4, dtype=np.uint32) if p.scheme == 1 else None
    with pytest.raises(mk.config_error):
        ctx.EvalCircuit(wires, [(0, 0, 2)], wb)
    for bad in ([(0, 1, 4)], [(0, 1, 2), (0, 1, 2)], [(0, 2, 3), (0, 1, 2)]):
        with pytest.raises(mk.MkfheError):
            ctx.EvalCircuit(wires, bad, wb)
    out, ob, levels = ctx.EvalCircuit(wires, np.zeros((0, 3), dtype=np.uint32), wb)  # empty netlist: nothing to do
    assert levels == 0 and (out == wires).all()
