"""Generate tests/golden/*.npz from the COMPILED REFERENCE (oracle/_ref/libmkfhe_ref.so, built in place from
/root/reference by `make -C oracle ref`).  Run in the CPU container:  python tests/golden/make_golden.py

Each fixture holds, for one toy parameter set: the parameters, the reference's own keys (bootstrapping key, Pkey, key-switch
key(s), ctNAND, secret material), and input/output pairs of every reference function on the hot path:
  ntt_in/ntt_fwd/ntt_inv           NativePoly::SetFormat
  dec_in/dec_out                   UniEncAccumulator::SignedDigitDecompose
  acc_in/acc_rot/acc_out           UniEncAccumulatorXZW{,_B}::EvalAcc
  g_ct1/g_b1/g_ct2/g_b2/g_out/g_bout/g_bits/g_m   BinFHEScheme::EvalBinGate(NAND) + Decrypt
Key generation in the reference is seeded from the clock, so regenerating changes the bytes (not the semantics).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import refapi  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

# name: (scheme, k, numberBits, cyclOrder, n, q, qKS, stdDev, baseKS, baseG, baseR)
TOYS = {
    "toy_mkntru_k2_N16": (0, 2, 27, 32, 4, 45181, 45181, 0.75, 32, 1 << 9, 32),
    "toy_mklwe_k2_N16": (1, 2, 27, 32, 4, 32749, 32749, 1.9, 32, 1 << 9, 2),
    "toy_mkntru_k3_N64_B7": (0, 3, 27, 128, 6, 45181, 45181, 0.75, 32, 1 << 7, 32),
    "toy_mklwe_k3_N32_B5": (1, 3, 27, 64, 5, 32749, 32749, 1.9, 32, 1 << 5, 2),
    "toy_mkntru_k4_N256": (0, 4, 27, 512, 8, 45181, 45181, 0.75, 32, 1 << 9, 32),
    "toy_mklwe_k2_N128": (1, 2, 27, 256, 7, 32749, 32749, 1.9, 32, 1 << 9, 2),
}


def make(name, args, seed):
    rng = np.random.default_rng(seed)
    r = refapi.Ref(*args)
    p = r.p
    r.keygen()
    d = dict(params=np.array([p.k, p.n, p.N, p.q, p.qKS, p.Q, p.baseG, p.digitsG, p.baseKS, p.dKS, p.psi, p.scheme,
                              args[2], args[3]], dtype=np.uint32))
    d["bk"], d["pkey"], d["sk"] = r.bk(), r.pkey(), r.sk()
    if p.scheme == 0:
        d["ksk"], bad = r.ksk_mntru()
        assert bad == 0
        d["ctnand"] = r.ctnand()
    else:
        d["kskA"], d["kskB"] = r.ksk_mklwe()
    x = rng.integers(0, p.Q, size=(3, p.N), dtype=np.uint32)
    x[2, :4] = [0, 1, p.Q - 1, p.Q >> 1]  # edge residues
    d["ntt_in"], d["ntt_fwd"], d["ntt_inv"] = x, r.ntt(x, False), r.ntt(x, True)
    d["dec_in"] = x
    d["dec_out"] = np.stack([r.decompose(row) for row in x])
    rot_mod = p.q if p.scheme == 0 else 2 * p.N
    acc_in = rng.integers(0, p.Q, size=(3, p.k * p.N), dtype=np.uint32)
    acc_rot = rng.integers(0, rot_mod, size=(3, p.k * p.n), dtype=np.uint32)
    acc_rot[1, :] = 0          # all-zero rotations (every step has the zero monomial)
    acc_rot[2, 0] = rot_mod - 1
    d["acc_in"], d["acc_rot"] = acc_in, acc_rot
    d["acc_out"] = np.stack([r.eval_acc(a, t, rot_mod) for a, t in zip(acc_in, acc_rot)])
    g = {k: [] for k in ("ct1", "b1", "ct2", "b2", "out", "bout", "bits", "m")}
    for m0, m1 in ((0, 0), (0, 1), (1, 0), (1, 1), (1, 0), (0, 0)):
        c1, b1 = r.encrypt(m0)
        c2, b2 = r.encrypt(m1)
        o, bo = r.eval_gate(c1, b1, c2, b2)
        for k_, v in zip(g, (c1, b1, c2, b2, o, bo, r.decrypt(o, bo), (m0, m1))):
            g[k_].append(v)
    for k_, v in g.items():
        d["g_" + k_] = np.array(v, dtype=np.uint32)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **d)
    print(name, {k_: v.shape for k_, v in d.items() if k_ in ("bk", "ksk", "kskA")}, "bits", d["g_bits"])


if __name__ == "__main__":
    for i, (name, args) in enumerate(TOYS.items()):
        make(name, args, 1000 + i)
