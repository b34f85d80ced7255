#!/bin/bash
# usage: tools/gpu_dev3.sh <tag> -- GPU parity tests, quick timings of the k = 2 / k = 4 shapes, NTT micro-benchmarks
TAG=$1; mkdir -p gpurun_out
(time python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_tests.log 2>&1; tail -3 gpurun_out/${TAG}_tests.log
L=gpurun_out/${TAG}_quick.log; : > $L
python tests/dev_quick_bench.py --paramset STD100_MKNTRU_LWE --batches 592,4096 --check 2 >> $L 2>&1
python tests/dev_quick_bench.py --paramset STD100_MKNTRU --batches 592,4096 --check 2 >> $L 2>&1
python tests/dev_quick_bench.py --paramset STD100_MKNTRU_2 --batches 444,1776 --check 1 >> $L 2>&1
grep -v "^$" $L | tail -20
for n in 1024 2048; do python bench.py --workload ntt --ntt-n $n --steps 5 --warmup 3 > gpurun_out/${TAG}_ntt$n.json 2> gpurun_out/${TAG}_ntt$n.err; cut -c1-1100 gpurun_out/${TAG}_ntt$n.json; tail -2 gpurun_out/${TAG}_ntt$n.err; done
