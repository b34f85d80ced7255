#!/bin/bash
# usage: tools/gpu_sanitize.sh <tag> -- compute-sanitizer memcheck + racecheck over smoke() (toy fixtures on the generic kernel,
# the specialised gate kernels on a short blind rotation) and over a 2-gate k = 4 run of the party-parallel kernel
TAG=$1; mkdir -p gpurun_out; L=gpurun_out/${TAG}_sanitizer.txt; : > $L
cat > /tmp/grouped_small.py <<'PY'
import numpy as np, mkfhe_b200 as mk
P = mk.derive_params(mk.PARAMSETS["STD100_MKNTRU_2"], mk.MKNTRU); P.n, P.device = 6, 0
rng = np.random.default_rng(1); ell = P.digitsG - 1
ctx = mk.BinFHEContext().GenerateFromParams(P)
ctx.LoadKeys(rng.integers(0, P.Q, size=(P.k * 2 * P.n + 1) * ell * 2 * P.N, dtype=np.uint32), rng.integers(0, P.Q, size=P.k * ell * P.N, dtype=np.uint32),
             ksk=rng.integers(0, P.qKS, size=P.k * P.N * P.dKS * P.n, dtype=np.uint32), ctnand=rng.integers(0, P.q, size=P.k * P.n, dtype=np.uint32))
out = ctx.EvalBinGate("NAND", rng.integers(0, P.q, size=(2, P.k, P.n), dtype=np.uint32), rng.integers(0, P.q, size=(2, P.k, P.n), dtype=np.uint32))
print("grouped small run done", out.shape); ctx.close()
PY
for tool in memcheck racecheck; do
  echo "==== compute-sanitizer --tool $tool : python __graft_entry__.py --smoke" >> $L
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python __graft_entry__.py --smoke 2>&1 | grep -v "^$" | tail -25 >> $L
  echo "==== compute-sanitizer --tool $tool : party-parallel kernel (MKFHE_GROUPS=2, k = 4, n = 6)" >> $L
  MKFHE_GROUPS=2 PYTHONPATH=. timeout 900 compute-sanitizer --tool $tool --print-limit 20 python /tmp/grouped_small.py 2>&1 | grep -v "^$" | tail -25 >> $L
done
grep -E "====|ERROR SUMMARY|RACECHECK SUMMARY|smoke|done" $L
