#!/bin/bash
# usage: tools/gpu_xzw.sh <tag> -- XZW (boolean-mkntru) timings at both register budgets + k = 4, with oracle checks
TAG=$1; mkdir -p gpurun_out; L=gpurun_out/${TAG}_quick.log; : > $L
q() { echo "== MKFHE_MINB=$1 $2" >> $L; MKFHE_MINB=$1 python tests/dev_quick_bench.py --paramset $2 --batches $3 --check $4 >> $L 2>&1; }
q 4 STD100_MKNTRU 592,4096 2
q 3 STD100_MKNTRU 444,3552 1
q 3 STD100_MKNTRU_2 444,1776 2
q 4 STD100_MKNTRU_LWE 592,4096 1
grep -v "^$" $L | tail -30
