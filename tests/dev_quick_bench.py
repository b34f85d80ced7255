#!/usr/bin/env python
"""Kernel-development loop: synthetic keys (uniform residues), dense random gate inputs, device-resident timing of the gate
kernel at several batch sizes, and a bit-exact check of a few gates against the C oracle (oracle/mkfhe_oracle.c).

    python tests/dev_quick_bench.py [--paramset STD100_MKNTRU_LWE] [--batches 592,4096] [--check 2] [--reps 3]
"""
import argparse, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)



def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--paramset", default="STD100_MKNTRU_LWE")
    ap.add_argument("--batches", default="592,4096")
    ap.add_argument("--check", type=int, default=2)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    import torch
    import mkfhe_b200 as mk
    from bench import make_workload
    os.environ["MKFHE_BENCH_SYNTHETIC_KEYS"] = "1"
    method = mk.MKNTRU_LWE if "LWE" in a.paramset else mk.MKNTRU
    batches = [int(x) for x in a.batches.split(",")]
    P, keys, ct1, b1, ct2, b2, _, _ = make_workload(a.paramset, method, max(batches), seed=7)
    ctx = mk.BinFHEContext().GenerateFromParams(P)
    lwe = ctx.is_lwe
    ctx.LoadKeys(keys["bk"], keys["pkey"], ksk=keys.get("ksk"), kskA=keys.get("kskA"), kskB=keys.get("kskB"), ctnand=keys.get("ctnand"))
    dev = torch.device("cuda")
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x).view(np.int32)).to(dev)
    for B in batches:
        d1, d2, db1, db2 = t(ct1[:B]), t(ct2[:B]), t(b1[:B]), t(b2[:B])
        out, bout = torch.empty_like(d1), torch.empty_like(db1)
        times = []
        for r in range(a.reps + 1):
            ctx.EvalBinGateDevice(d1, d2, out, db1 if lwe else None, db2 if lwe else None, bout if lwe else None)
            rot, ks = ctx.last_gate_timing()
            times.append(rot + ks)
        best = min(times[1:])
        print(f"{a.paramset} batch {B:6d}: {best:9.3f} ms  ({B / best * 1e3:9.1f} gates/s)  all: {' '.join(f'{x:.2f}' for x in times)}", flush=True)
    if a.check:
        from oracle import portapi
        o = portapi.Oracle(k=P.k, N=P.N, Q=P.Q, psi=P.psi, q=P.q, qKS=P.qKS, n=P.n, baseG_log2=P.baseG_log2, digitsG=P.digitsG,
                           baseKS=P.baseKS, dKS=P.dKS, variant=P.variant, scheme=P.scheme)
        B = batches[-1]
        got = out.cpu().numpy().view(np.uint32)
        gb = bout.cpu().numpy().view(np.uint32)
        t0 = time.time()
        for i in list(range(a.check - 1)) + [B - 1]:
            eo, eb = o.eval_nand(keys["bk"], keys["pkey"], keys.get("kskA", keys.get("ksk")), keys.get("kskB"), keys.get("ctnand"),
                                 ct1[i].ravel(), int(b1[i]), ct2[i].ravel(), int(b2[i]))
            ok = (got[i].ravel() == eo).all() and (not lwe or int(gb[i]) == eb)
            print(f"gate {i}: {'bit-exact vs oracle' if ok else 'MISMATCH'}", flush=True)
            if not ok:
                sys.exit(1)
        print(f"oracle check {time.time() - t0:.1f}s")
    ctx.close()


if __name__ == "__main__":
    main()
