#!/bin/bash
# usage: tools/gpu_minb.sh <tag> -- register-budget sweep of the gate kernel (MKFHE_MINB = gates promised per SM)
TAG=$1; mkdir -p gpurun_out; L=gpurun_out/${TAG}_quick.log; : > $L
q() { echo "== MKFHE_MINB=$1 $2" >> $L; MKFHE_MINB=$1 python tests/dev_quick_bench.py --paramset $2 --batches $3 --check $4 >> $L 2>&1; }
q 4 STD100_MKNTRU_LWE 592,4096 0
q 3 STD100_MKNTRU_LWE 444,3552 2
q 4 STD100_MKNTRU 592,4096 0
q 3 STD100_MKNTRU 444,3552 2
q 4 STD100_MKNTRU_2 444,1776 0
q 3 STD100_MKNTRU_2 444,1776 2
q 4 STD100_MKNTRU_3 296,1184 0
q 2 STD100_MKNTRU_3 296,1184 1
grep -v "^$" $L | tail -40
