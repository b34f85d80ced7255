#!/bin/bash
# usage: tools/gpu_sweep.sh <tag> -- one GPU box visit: GPU tests, config 3 party sweep, config 4 (1 GPU, reduced), config 5
TAG=$1; mkdir -p gpurun_out
(time python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_tests.log 2>&1; tail -3 gpurun_out/${TAG}_tests.log
run() { name=$1; shift; (time timeout 900 python bench.py "$@") > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err; \
        echo "== $name"; cut -c1-700 gpurun_out/${TAG}_${name}.json; tail -2 gpurun_out/${TAG}_${name}.err; }
run ntt2048 --workload ntt --ntt-n 2048 --steps 5 --warmup 3
run ntt1024 --workload ntt --ntt-n 1024 --steps 5 --warmup 3
run k2  --paramset STD100_MKNTRU   --batch 16384 --steps 1 --warmup 1 --synthetic-keys --check-oracle 2 --no-cpu-baseline
run k4  --paramset STD100_MKNTRU_2 --batch 16384 --steps 1 --warmup 1 --synthetic-keys --check-oracle 2 --no-cpu-baseline
run k8  --paramset STD100_MKNTRU_3 --batch 4096  --steps 1 --warmup 1 --synthetic-keys --check-oracle 1 --no-cpu-baseline
run k16 --paramset STD100_MKNTRU_4 --batch 1184  --steps 1 --warmup 1 --synthetic-keys --check-oracle 1 --no-cpu-baseline
run sharded_k4 --workload sharded --paramset STD100_MKNTRU_2 --total-gates 32768 --steps 1 --warmup 1 --check-oracle 1
