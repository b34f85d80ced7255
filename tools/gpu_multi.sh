#!/bin/bash
# usage: tools/gpu_multi.sh <tag> <ngpus> -- multi-GPU visit: default bench under torchrun, sharded job, reference arm
TAG=$1; N=$2; mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
(time timeout 900 $TR bench.py --gpus $N --steps 3 --warmup 3) > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
cut -c1-900 gpurun_out/${TAG}_bench_n$N.json; tail -3 gpurun_out/${TAG}_bench_n$N.err
(time timeout 1500 $TR bench.py --gpus $N --workload sharded --paramset STD100_MKNTRU_2 --total-gates ${TOTAL:-32768} --steps 1 --warmup 1) \
   > gpurun_out/${TAG}_sharded_n$N.json 2> gpurun_out/${TAG}_sharded_n$N.err
cut -c1-1200 gpurun_out/${TAG}_sharded_n$N.json; tail -3 gpurun_out/${TAG}_sharded_n$N.err
(time timeout 600 $TR bench.py --gpus $N --impl reference --steps 1 --warmup 1) > gpurun_out/${TAG}_ref_n$N.json 2> gpurun_out/${TAG}_ref_n$N.err
cut -c1-600 gpurun_out/${TAG}_ref_n$N.json; tail -3 gpurun_out/${TAG}_ref_n$N.err
python bench.py --workload ntt --ntt-n 2048 --steps 5 --warmup 3 > gpurun_out/${TAG}_ntt2048.json 2> gpurun_out/${TAG}_ntt2048.err
cut -c1-900 gpurun_out/${TAG}_ntt2048.json; tail -2 gpurun_out/${TAG}_ntt2048.err
