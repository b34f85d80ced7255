"""Shared helpers for the test-suite (fixtures -> oracle / GPU contexts)."""
import glob
import os
from types import SimpleNamespace

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def golden_names():
    return sorted(os.path.basename(f)[:-4] for f in glob.glob(os.path.join(GOLDEN, "*.npz")))


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    d = {k: z[k] for k in z.files}
    v = [int(x) for x in d["params"]]
    p = SimpleNamespace(k=v[0], n=v[1], N=v[2], q=v[3], qKS=v[4], Q=v[5], baseG=v[6], digitsG=v[7], baseKS=v[8], dKS=v[9],
                        psi=v[10], scheme=v[11], numberBits=v[12], cyclOrder=v[13])
    p.ell = p.digitsG - 1
    p.variant = p.scheme          # MKNTRU -> XZW (0), MKNTRU_LWE -> XZW_B (1)
    p.baseG_log2 = p.baseG.bit_length() - 1
    p.rot_mod = p.q if p.scheme == 0 else 2 * p.N
    d["p"] = p
    d["kskA_any"] = d["ksk"] if p.scheme == 0 else d["kskA"]
    d["kskB_any"] = None if p.scheme == 0 else d["kskB"]
    d["ctnand_any"] = d.get("ctnand")
    return d


def make_oracle(p):
    from oracle import portapi

    return portapi.Oracle(k=p.k, N=p.N, Q=p.Q, psi=p.psi, q=p.q, qKS=p.qKS, n=p.n, baseG_log2=p.baseG_log2,
                          digitsG=p.digitsG, baseKS=p.baseKS, dKS=p.dKS, variant=p.variant, scheme=p.scheme)


def params_from_ref(rp):
    """RefParams (oracle/refapi.py) -> the namespace used above."""
    p = SimpleNamespace(k=rp.k, n=rp.n, N=rp.N, q=rp.q, qKS=rp.qKS, Q=rp.Q, baseG=rp.baseG, digitsG=rp.digitsG,
                        baseKS=rp.baseKS, dKS=rp.dKS, psi=rp.psi, scheme=rp.scheme)
    p.ell = p.digitsG - 1
    p.variant = p.scheme
    p.baseG_log2 = p.baseG.bit_length() - 1
    p.rot_mod = p.q if p.scheme == 0 else 2 * p.N
    return p
