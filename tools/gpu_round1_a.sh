#!/bin/bash
# GPU session A: pipe microbench, generic-path bench on dense inputs, ncu launch list + full capture of the generic kernel
mkdir -p gpurun_out
tools/_build/pipe_microbench > gpurun_out/r1_pipe_microbench.json 2>&1
cat gpurun_out/r1_pipe_microbench.json
(time python bench.py --steps 3 --warmup 3) > gpurun_out/r1_bench_generic_dense.json 2> gpurun_out/r1_bench_generic_dense.err
cat gpurun_out/r1_bench_generic_dense.json; tail -3 gpurun_out/r1_bench_generic_dense.err
CMD="python bench.py --batch 296 --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/r1_ncu_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1_generic_launches.csv $CMD > gpurun_out/r1_ncu_launches.log 2>&1
tail -2 gpurun_out/r1_ncu_plain.log | cut -c1-400
$CMD > gpurun_out/r1_ncu_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:blind_rotate -s 4 -c 1 -o gpurun_out/r1_generic_blind_rotate $CMD > gpurun_out/r1_ncu_full.log 2>&1
tail -3 gpurun_out/r1_ncu_full.log
ls -la gpurun_out
