#!/bin/bash
# usage: tools/gpu_dev.sh <tag> -- kernel-development visit: GPU parity tests, then tests/dev_quick_bench.py on the k = 2 shapes and k = 4
TAG=$1; mkdir -p gpurun_out
(time python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_tests.log 2>&1; tail -3 gpurun_out/${TAG}_tests.log
L=gpurun_out/${TAG}_quick.log; : > $L
python tests/dev_quick_bench.py --paramset STD100_MKNTRU_LWE --batches 592,4096 --check 2 >> $L 2>&1
python tests/dev_quick_bench.py --paramset STD100_MKNTRU --batches 592,4096 --check 2 >> $L 2>&1
MKFHE_FAST_ALIAS=0 python tests/dev_quick_bench.py --paramset STD100_MKNTRU_2 --batches 444,1776 --check 1 >> $L 2>&1
MKFHE_FAST_ALIAS=1 python tests/dev_quick_bench.py --paramset STD100_MKNTRU_2 --batches 592,2368 --check 2 >> $L 2>&1
python tests/dev_quick_bench.py --paramset STD100_MKNTRU_3 --batches 296,1184 --check 1 >> $L 2>&1
grep -v "^$" $L | tail -40
