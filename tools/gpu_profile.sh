#!/bin/bash
# usage: tools/gpu_profile.sh <tag> [bench args...]   -- plain run, ncu launch list, ncu --set full of the blind-rotation kernel
TAG=$1; shift
mkdir -p gpurun_out
CMD="python bench.py --batch 592 --steps 1 --warmup 3 --no-cpu-baseline $*"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_launches.log 2>&1
cut -c1-300 gpurun_out/${TAG}_plain.json
$CMD > gpurun_out/${TAG}_plain2.json 2> gpurun_out/${TAG}_plain2.err && \
ncu --set full --clock-control none --import-source on -k regex:blind_rotate -s 5 -c 1 -o gpurun_out/${TAG}_blind_rotate $CMD > gpurun_out/${TAG}_full.log 2>&1
tail -2 gpurun_out/${TAG}_full.log
