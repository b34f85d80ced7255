#!/bin/bash
# usage: tools/gpu_prof4.sh <tag> -- ncu --set full of the one-group gate kernel in its 128-register build (296 gates, boolean-mklwe)
TAG=$1; mkdir -p gpurun_out; name=mklwe296
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --synthetic-keys --batch 296"
$CMD > gpurun_out/${TAG}_${name}_plain.json 2> gpurun_out/${TAG}_${name}_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:blind_rotate -s 3 -c 1 -o gpurun_out/${TAG}_${name} $CMD > gpurun_out/${TAG}_${name}_full.log 2>&1
python tools/ncu_summary.py gpurun_out/${TAG}_${name}.ncu-rep > gpurun_out/${TAG}_${name}.txt 2>&1
python tools/ncu_stalls.py gpurun_out/${TAG}_${name}.ncu-rep >> gpurun_out/${TAG}_${name}.txt 2>&1
rm -f gpurun_out/${TAG}_${name}.ncu-rep; head -60 gpurun_out/${TAG}_${name}.txt
