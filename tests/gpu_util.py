"""GPU-side helpers for the parity tests: build a mkfhe_b200 context from a fixture / reference context."""
import numpy as np

import mkfhe_b200 as mk
from mkfhe_b200.lib import Params


def ctx_from_params(p, device=-1):
    """p: namespace from tests/util.py (load_golden / params_from_ref)."""
    P = Params(k=p.k, N=p.N, Q=p.Q, psi=p.psi, q=p.q, qKS=p.qKS, n=p.n, baseG_log2=p.baseG_log2, digitsG=p.digitsG,
               baseKS=p.baseKS, dKS=p.dKS, variant=p.variant, scheme=p.scheme, device=device)
    return mk.BinFHEContext().GenerateFromParams(P)


def load_fixture_keys(ctx, g):
    p = g["p"]
    if p.scheme == 0:
        ctx.LoadKeys(g["bk"], g["pkey"], ksk=g["ksk"], ctnand=g["ctnand"])
    else:
        ctx.LoadKeys(g["bk"], g["pkey"], kskA=g["kskA"], kskB=g["kskB"])


def load_ref_keys(ctx, r):
    """r: oracle.refapi.Ref after keygen.  Returns the exported arrays for reuse by the C oracle."""
    keys = dict(bk=r.bk(), pkey=r.pkey(), sk=r.sk())
    if r.p.scheme == 0:
        keys["kskA"], bad = r.ksk_mntru()
        assert bad == 0
        keys["kskB"], keys["ctnand"] = None, r.ctnand()
        ctx.LoadKeys(keys["bk"], keys["pkey"], ksk=keys["kskA"], ctnand=keys["ctnand"])
    else:
        keys["kskA"], keys["kskB"] = r.ksk_mklwe()
        keys["ctnand"] = None
        ctx.LoadKeys(keys["bk"], keys["pkey"], kskA=keys["kskA"], kskB=keys["kskB"])
    return keys
