#!/bin/bash
# usage: tools/gpu_k16.sh <tag> -- k = 16 (STD100_MKNTRU_4): one group / two groups / four groups per gate, oracle-checked
TAG=$1; mkdir -p gpurun_out; L=gpurun_out/${TAG}_quick.log; : > $L
q() { echo "== MKFHE_GROUPS=$1 $2" >> $L; MKFHE_GROUPS=$1 python tests/dev_quick_bench.py --paramset $2 --batches $3 --check $4 --reps 2 >> $L 2>&1; }
q 1 STD100_MKNTRU_4 148,296 0
q 2 STD100_MKNTRU_4 148,296 1
q 4 STD100_MKNTRU_4 148,296 1
q 2 STD100_MKNTRU_3 148,592 1
grep -v "^$" $L | tail -30
(MKFHE_GROUPS=2 timeout 600 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -k "gate_batch or eval_acc") 2>&1 | tail -3
