#!/bin/bash
# usage: tools/gpu_dev2.sh <tag> -- party-parallel kernel: parity tests with groups forced, then timings per party count
TAG=$1; mkdir -p gpurun_out
(MKFHE_GROUPS=2 timeout 600 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q) > gpurun_out/${TAG}_tests_g2.log 2>&1; tail -3 gpurun_out/${TAG}_tests_g2.log
L=gpurun_out/${TAG}_quick.log; : > $L
q() { echo "== MKFHE_GROUPS=$1 $2" >> $L; MKFHE_GROUPS=$1 timeout 900 python tests/dev_quick_bench.py --paramset $2 --batches $3 --check $4 >> $L 2>&1; }
q 2 STD100_MKNTRU_2 296,1184 2
q 1 STD100_MKNTRU_3 296,1184 0
q 4 STD100_MKNTRU_3 148,592 2
q 2 STD100_MKNTRU_3 148,592 1
q 1 STD100_MKNTRU_4 148,296 0
q 4 STD100_MKNTRU_4 148,296 1
q 2 STD100_MKNTRU_LWE 296,2368 2
grep -v "^$" $L | tail -60
