"""GPU: the larger party counts of BASELINE.json config 3 / 4 (STD100_MKNTRU_2/_3/_4: k = 4, 8, 16) against the C oracle, on
uniform random keys and dense random ciphertexts -- every kernel variant the host can select: 80- and 128-register budgets
(k = 4, 8), the party-parallel kernel (k = 16 by default; forced for k = 4 with MKFHE_GROUPS), the aliased exchange buffer."""
import os

import numpy as np
import pytest

import mkfhe_b200 as mk
from concurrent.futures import ThreadPoolExecutor
from oracle import refapi
from util import make_oracle

pytestmark = pytest.mark.gpu

# gates re-computed on the CPU oracle per party count (16 for k <= 4, 4 for k = 8, 1 for k = 16) on the
# kernel the host picks by default; the forced variants check fewer (the oracle costs 0.3 s x 3 k (k + 1) / 18 per gate)
FULL = {2: 16, 4: 16, 8: 4, 16: 1}

CASES = [  # paramset, environment overrides, gates in the batch, gates re-computed on the CPU oracle
    # batches up to the SM count run on the party-parallel kernel by default (two groups); MKFHE_GROUPS=1 selects the
    # one-group kernel that larger batches use
    ("STD100_MKNTRU_LWE", {}, 17, 16),
    ("STD100_MKNTRU_LWE", {"MKFHE_GROUPS": "1"}, 17, 16),
    ("STD100_MKNTRU", {}, 17, 16),
    ("STD100_MKNTRU", {"MKFHE_GROUPS": "1"}, 17, 16),
    ("STD100_MKNTRU", {"MKFHE_GROUPS": "1", "MKFHE_MINB": "4"}, 5, 4),  # the 64-register XZW build (batches that fill 4-gate waves better)
    ("STD100_MKNTRU", {"MKFHE_GROUPS": "3"}, 3, 2),
    ("STD100_MKNTRU_LWE_2", {"MKFHE_GROUPS": "3"}, 3, 2),
    ("STD100_MKNTRU_2", {"MKFHE_GROUPS": "1"}, 17, 16),
    ("STD100_MKNTRU_2", {}, 17, 16),
    ("STD100_MKNTRU_2", {"MKFHE_GROUPS": "4"}, 3, 2),
    ("STD100_MKNTRU_2", {"MKFHE_GROUPS": "1", "MKFHE_FAST_ALIAS": "1"}, 3, 2),
    ("STD100_MKNTRU_LWE_2", {"MKFHE_GROUPS": "1", "MKFHE_MINB": "4"}, 3, 2),
    ("STD100_MKNTRU_LWE_2", {}, 5, 4),
    ("STD100_MKNTRU_3", {"MKFHE_GROUPS": "1"}, 5, 4),
    ("STD100_MKNTRU_3", {}, 5, 4),
    ("STD100_MKNTRU_4", {}, 2, 1),
    # thread-block clusters (blind_rotate_cluster): 2 / 4 / 8 SMs per gate, forced; batches of at most SMs / size gates pick the
    # largest size that fits by default (the 17-, 5- and 2-gate default cases above), MKFHE_CLUSTER=0 keeps them on one SM
    ("STD100_MKNTRU_LWE", {"MKFHE_CLUSTER": "2"}, 5, 4),
    ("STD100_MKNTRU", {"MKFHE_CLUSTER": "2"}, 5, 4),
    ("STD100_MKNTRU_LWE", {"MKFHE_CLUSTER": "0"}, 17, 4),
    ("STD100_MKNTRU", {"MKFHE_CLUSTER": "0"}, 17, 4),
    ("STD100_MKNTRU_2", {"MKFHE_CLUSTER": "2"}, 3, 2),
    ("STD100_MKNTRU_2", {"MKFHE_CLUSTER": "4"}, 3, 2),
    ("STD100_MKNTRU_LWE_2", {"MKFHE_CLUSTER": "4"}, 3, 2),
    ("STD100_MKNTRU_2", {"MKFHE_CLUSTER": "0"}, 3, 2),
    ("STD100_MKNTRU_3", {"MKFHE_CLUSTER": "8"}, 2, 1),
    ("STD100_MKNTRU_LWE_3", {"MKFHE_CLUSTER": "4"}, 2, 1),
    ("STD100_MKNTRU_4", {"MKFHE_CLUSTER": "8"}, 2, 1),
    ("STD100_MKNTRU_4", {"MKFHE_CLUSTER": "0"}, 2, 1),
    # accumulators in global memory (blind_rotate_fast_ag, an experiment behind MKFHE_AG=1; batches above the SM count only)
    ("STD100_MKNTRU_LWE", {"MKFHE_AG": "1"}, 150, 2),
    ("STD100_MKNTRU_2", {"MKFHE_AG": "1", "MKFHE_MINB": "4"}, 150, 2),
    ("STD100_MKNTRU_4", {"MKFHE_GROUPS": "1"}, 2, 1),
]


@pytest.mark.parametrize("paramset,env,gates,checked", CASES, ids=[f"{c[0]}-{'-'.join(k + v for k, v in c[1].items()) or 'default'}" for c in CASES])
def test_large_party_counts_vs_oracle(paramset, env, gates, checked, monkeypatch):
    if paramset not in mk.PARAMSETS:
        pytest.skip(f"{paramset} not in the parameter table")
    for k_, v in env.items():
        monkeypatch.setenv(k_, v)
    method = mk.MKNTRU_LWE if "LWE" in paramset else mk.MKNTRU
    P = mk.derive_params(mk.PARAMSETS[paramset], method)
    rng = np.random.default_rng(31 + P.k)
    ell, S, lwe = P.digitsG - 1, (2 if P.variant == 0 else 1), P.scheme == 1
    bk = rng.integers(0, P.Q, size=(P.k * S * P.n + 1) * ell * 2 * P.N, dtype=np.uint32)
    pkey = rng.integers(0, P.Q, size=P.k * ell * P.N, dtype=np.uint32)
    ctx = mk.BinFHEContext().GenerateFromParams(P)
    if lwe:
        rows = P.k * P.N * P.baseKS * P.dKS
        kskA = rng.integers(0, P.qKS, size=rows * P.n, dtype=np.uint32)
        kskB = rng.integers(0, P.qKS, size=rows, dtype=np.uint32)
        ctnand = None
        ctx.LoadKeys(bk, pkey, kskA=kskA, kskB=kskB)
    else:
        kskA = rng.integers(0, P.qKS, size=P.k * P.N * P.dKS * P.n, dtype=np.uint32)
        kskB = None
        ctnand = rng.integers(0, P.q, size=P.k * P.n, dtype=np.uint32)
        ctx.LoadKeys(bk, pkey, ksk=kskA, ctnand=ctnand)
    ct1 = rng.integers(0, P.q, size=(gates, P.k, P.n), dtype=np.uint32)
    ct2 = rng.integers(0, P.q, size=(gates, P.k, P.n), dtype=np.uint32)
    b1 = rng.integers(0, P.q, size=gates, dtype=np.uint32)
    b2 = rng.integers(0, P.q, size=gates, dtype=np.uint32)
    pick, dup = 0, gates - 1  # a copy of a checked gate elsewhere in the batch must come out identical
    ct1[dup], ct2[dup], b1[dup], b2[dup] = ct1[pick], ct2[pick], b1[pick], b2[pick]
    if lwe:
        out, bout = ctx.EvalBinGate("NAND", ct1, ct2, b1, b2)
    else:
        out, bout = ctx.EvalBinGate("NAND", ct1, ct2), np.zeros(gates, dtype=np.uint32)
    out = np.asarray(out)
    o = make_oracle(P)

    def one(i):
        eo, eb = o.eval_nand(bk, pkey, kskA, kskB, ctnand, ct1[i].ravel(), int(b1[i]), ct2[i].ravel(), int(b2[i]))
        return bool((out[i].ravel() == eo).all() and (not lwe or int(bout[i]) == eb))

    with ThreadPoolExecutor(max_workers=os.cpu_count() or 1) as ex:  # the C oracle releases the GIL
        ok = list(ex.map(one, range(checked)))
    assert all(ok), f"gates {[i for i, v in enumerate(ok) if not v]} differ from the oracle"
    assert (out[dup] == out[pick]).all() and int(bout[dup]) == int(bout[pick])
    ctx.close()


@pytest.mark.skipif(not refapi.available(), reason="oracle/_ref/libmkfhe_ref.so not present")
def test_k4_full_size_against_the_compiled_reference():
    """the party sweep never touched the unmodified reference above k = 2.  One full-size k = 4 context
    (STD100_MKNTRU_2: the reference's own keygen, 140 MiB bootstrapping key): EvalAcc on a random accumulator and one NAND
    gate on the reference's encryptions, bit for bit against UniEncAccumulatorXZW::EvalAcc / BinFHEScheme::EvalBinGate."""
    from gpu_util import ctx_from_params, load_ref_keys
    from util import params_from_ref

    r = refapi.Ref.from_paramset("STD100_MKNTRU_2")
    r.keygen()
    p = params_from_ref(r.p)
    assert p.k == 4
    ctx = ctx_from_params(p)
    load_ref_keys(ctx, r)
    rng = np.random.default_rng(44)
    acc = rng.integers(0, p.Q, size=(1, p.k, p.N), dtype=np.uint32)
    rot = rng.integers(0, p.rot_mod, size=(1, p.k, p.n), dtype=np.uint32)
    got = ctx.EvalAcc(acc, rot, p.rot_mod)
    assert (got[0].ravel() == r.eval_acc(acc[0].ravel(), rot[0].ravel(), p.rot_mod)).all()
    c1, _ = r.encrypt(1)
    c2, _ = r.encrypt(0)
    eo, _ = r.eval_gate(c1, 0, c2, 0)
    for B in (1, 149):  # the party-parallel (latency) kernel and the one-group kernel
        out = np.asarray(ctx.EvalBinGate("NAND", np.tile(c1.reshape(1, p.k, p.n), (B, 1, 1)), np.tile(c2.reshape(1, p.k, p.n), (B, 1, 1))))
        assert (out.reshape(B, -1) == eo).all()
    assert r.decrypt(eo, 0) == 1
    ctx.close()
    r.close()


@pytest.mark.parametrize("k,cl", [(3, "2"), (5, "4"), (6, "4")])
def test_cluster_kernel_party_counts_that_do_not_divide(k, cl, monkeypatch):
    """blind_rotate_cluster with a party count that is not a multiple of the cluster size (CTAs own different numbers of parties;
    the reference's tables stop at powers of two, its code does not): one gate and one EvalAcc against the C oracle, and against
    the one-SM kernels."""
    P = mk.derive_params(mk.PARAMSETS["STD100_MKNTRU"], mk.MKNTRU)
    P.k = k
    rng = np.random.default_rng(100 + k)
    ell, S = P.digitsG - 1, 2
    bk = rng.integers(0, P.Q, size=(P.k * S * P.n + 1) * ell * 2 * P.N, dtype=np.uint32)
    pkey = rng.integers(0, P.Q, size=P.k * ell * P.N, dtype=np.uint32)
    ksk = rng.integers(0, P.qKS, size=P.k * P.N * P.dKS * P.n, dtype=np.uint32)
    ctnand = rng.integers(0, P.q, size=P.k * P.n, dtype=np.uint32)
    ct1 = rng.integers(0, P.q, size=(2, P.k, P.n), dtype=np.uint32)
    ct2 = rng.integers(0, P.q, size=(2, P.k, P.n), dtype=np.uint32)
    acc = rng.integers(0, P.Q, size=(1, P.k, P.N), dtype=np.uint32)
    rot = rng.integers(0, P.q, size=(1, P.k, P.n), dtype=np.uint32)
    res = {}
    for mode in (cl, "0"):
        monkeypatch.setenv("MKFHE_CLUSTER", mode)
        ctx = mk.BinFHEContext().GenerateFromParams(P)
        ctx.LoadKeys(bk, pkey, ksk=ksk, ctnand=ctnand)
        res[mode] = (np.asarray(ctx.EvalBinGate("NAND", ct1, ct2)), np.asarray(ctx.EvalAcc(acc, rot, P.q)))
        ctx.close()
    assert (res[cl][0] == res["0"][0]).all() and (res[cl][1] == res["0"][1]).all()
    o = make_oracle(P)
    eo, _ = o.eval_nand(bk, pkey, ksk, None, ctnand, ct1[1].ravel(), 0, ct2[1].ravel(), 0)
    assert (res[cl][0][1].ravel() == eo).all()
    ea = o.eval_acc(bk, pkey, acc[0].ravel(), rot[0].ravel(), P.q)
    assert (res[cl][1][0].ravel() == np.asarray(ea).ravel()).all()


def test_cluster_policy_reports_sms_per_gate(monkeypatch):
    """mkfhe_last_gate_sms: batches of which the GPU holds one cluster per gate at once run on 2 (k = 2) or 4 (k = 4) SMs per gate,
    larger ones on one SM; MKFHE_CLUSTER=0 turns the clusters off.  Outputs do not depend on the choice."""
    import torch

    sms = torch.cuda.get_device_properties(0).multi_processor_count
    for paramset, want in (("STD100_MKNTRU_LWE", 2), ("STD100_MKNTRU_2", 4)):
        monkeypatch.delenv("MKFHE_CLUSTER", raising=False)
        method = mk.MKNTRU_LWE if "LWE" in paramset else mk.MKNTRU
        P = mk.derive_params(mk.PARAMSETS[paramset], method)
        P.n = 16  # a short blind rotation: the policy does not depend on n
        rng = np.random.default_rng(9)
        ell, S, lwe = P.digitsG - 1, (2 if P.variant == 0 else 1), P.scheme == 1
        bk = rng.integers(0, P.Q, size=(P.k * S * P.n + 1) * ell * 2 * P.N, dtype=np.uint32)
        pkey = rng.integers(0, P.Q, size=P.k * ell * P.N, dtype=np.uint32)
        ctx = mk.BinFHEContext().GenerateFromParams(P)
        if lwe:
            rows = P.k * P.N * P.baseKS * P.dKS
            ctx.LoadKeys(bk, pkey, kskA=rng.integers(0, P.qKS, size=rows * P.n, dtype=np.uint32),
                         kskB=rng.integers(0, P.qKS, size=rows, dtype=np.uint32))
        else:
            ctx.LoadKeys(bk, pkey, ksk=rng.integers(0, P.qKS, size=P.k * P.N * P.dKS * P.n, dtype=np.uint32),
                         ctnand=rng.integers(0, P.q, size=P.k * P.n, dtype=np.uint32))
        B = sms + 1
        ct1 = rng.integers(0, P.q, size=(B, P.k, P.n), dtype=np.uint32)
        ct2 = rng.integers(0, P.q, size=(B, P.k, P.n), dtype=np.uint32)
        b1 = rng.integers(0, P.q, size=B, dtype=np.uint32)
        b2 = rng.integers(0, P.q, size=B, dtype=np.uint32)
        run = (lambda m: ctx.EvalBinGate("NAND", ct1[:m], ct2[:m], b1[:m], b2[:m])[0]) if lwe else (lambda m: ctx.EvalBinGate("NAND", ct1[:m], ct2[:m]))
        big = np.asarray(run(B))
        assert ctx.last_gate_sms() == 1
        small = np.asarray(run(3))
        assert ctx.last_gate_sms() == want
        assert (small == big[:3]).all()
        monkeypatch.setenv("MKFHE_CLUSTER", "0")
        off = np.asarray(run(3))
        assert ctx.last_gate_sms() == 1 and (off == small).all()
        ctx.close()
