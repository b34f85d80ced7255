#!/bin/bash
# usage: tools/gpu_sweep2.sh <tag> -- one GPU: full GPU test suite, default bench, config 3 party sweep at batch 16384, config 5,
# key-generation bench
TAG=$1; mkdir -p gpurun_out
(time python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_tests.log 2>&1; head -3 gpurun_out/${TAG}_tests.log
run() { name=$1; shift; (time timeout 1500 python bench.py "$@") > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err; \
        echo "== $name"; cut -c1-420 gpurun_out/${TAG}_${name}.json; tail -1 gpurun_out/${TAG}_${name}.err; }
run bench --steps 3 --warmup 3
run keygen_mklwe --workload keygen --paramset STD100_MKNTRU_LWE --steps 5 --warmup 3
run keygen_mkntru --workload keygen --paramset STD100_MKNTRU --steps 5 --warmup 3
run ntt2048 --workload ntt --ntt-n 2048 --steps 5 --warmup 3
run ntt1024 --workload ntt --ntt-n 1024 --steps 5 --warmup 3
run k2  --paramset STD100_MKNTRU   --batch 16384 --steps 1 --warmup 1 --synthetic-keys --check-oracle 2 --no-cpu-baseline
run k4  --paramset STD100_MKNTRU_2 --batch 16384 --steps 1 --warmup 1 --synthetic-keys --check-oracle 2 --no-cpu-baseline
run k8  --paramset STD100_MKNTRU_3 --batch 16384 --steps 1 --warmup 1 --synthetic-keys --check-oracle 1 --no-cpu-baseline
run k16 --paramset STD100_MKNTRU_4 --batch 16384 --steps 1 --warmup 1 --synthetic-keys --check-oracle 1 --no-cpu-baseline
