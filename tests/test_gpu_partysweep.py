"""GPU: the larger party counts of BASELINE.json config 3 / 4 (STD100_MKNTRU_2/_3/_4: k = 4, 8, 16) against the C oracle, on
uniform random keys and dense random ciphertexts -- every kernel variant the host can select: 80- and 128-register budgets
(k = 4, 8), the party-parallel kernel (k = 16 by default; forced for k = 4 with MKFHE_GROUPS), the aliased exchange buffer."""
import os

import numpy as np
import pytest

import mkfhe_b200 as mk
from util import make_oracle

pytestmark = pytest.mark.gpu

CASES = [  # paramset, environment overrides, gates, which gate to re-compute on the CPU oracle
    # batches up to the SM count run on the party-parallel kernel by default (two groups); MKFHE_GROUPS=1 selects the
    # one-group kernel that larger batches use
    ("STD100_MKNTRU_LWE", {}, 3, 2),
    ("STD100_MKNTRU_LWE", {"MKFHE_GROUPS": "1"}, 3, 0),
    ("STD100_MKNTRU", {}, 3, 1),
    ("STD100_MKNTRU", {"MKFHE_GROUPS": "1"}, 3, 0),
    ("STD100_MKNTRU_2", {"MKFHE_GROUPS": "1"}, 5, 4),
    ("STD100_MKNTRU_2", {}, 3, 0),
    ("STD100_MKNTRU_2", {"MKFHE_GROUPS": "4"}, 3, 1),
    ("STD100_MKNTRU_2", {"MKFHE_GROUPS": "1", "MKFHE_FAST_ALIAS": "1"}, 3, 2),
    ("STD100_MKNTRU_LWE_2", {"MKFHE_GROUPS": "1", "MKFHE_MINB": "4"}, 3, 1),
    ("STD100_MKNTRU_LWE_2", {}, 3, 2),
    ("STD100_MKNTRU_3", {"MKFHE_GROUPS": "1"}, 2, 1),
    ("STD100_MKNTRU_3", {}, 2, 0),
    ("STD100_MKNTRU_4", {}, 2, 0),
]


@pytest.mark.parametrize("paramset,env,gates,pick", CASES, ids=[f"{c[0]}-{'-'.join(k + v for k, v in c[1].items()) or 'default'}" for c in CASES])
def test_large_party_counts_vs_oracle(paramset, env, gates, pick, monkeypatch):
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
    dup = (pick + 1) % gates  # a copy of the checked gate elsewhere in the batch must come out identical
    ct1[dup], ct2[dup], b1[dup], b2[dup] = ct1[pick], ct2[pick], b1[pick], b2[pick]
    if lwe:
        out, bout = ctx.EvalBinGate("NAND", ct1, ct2, b1, b2)
    else:
        out, bout = ctx.EvalBinGate("NAND", ct1, ct2), np.zeros(gates, dtype=np.uint32)
    o = make_oracle(P)
    eo, eb = o.eval_nand(bk, pkey, kskA, kskB, ctnand, ct1[pick].ravel(), int(b1[pick]), ct2[pick].ravel(), int(b2[pick]))
    assert (np.asarray(out)[pick].ravel() == eo).all() and (not lwe or int(bout[pick]) == eb)
    assert (np.asarray(out)[dup] == np.asarray(out)[pick]).all() and int(bout[dup]) == int(bout[pick])
    ctx.close()
