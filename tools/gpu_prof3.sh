#!/bin/bash
# usage: tools/gpu_prof3.sh <tag> -- ncu --set full of the party-parallel gate kernel in latency mode (148 gates, k = 2, both variants)
TAG=$1; mkdir -p gpurun_out
prof() { name=$1; shift; CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --synthetic-keys $*"
  $CMD > gpurun_out/${TAG}_${name}_plain.json 2> gpurun_out/${TAG}_${name}_plain.err || return
  ncu --set full --clock-control none --import-source on -k regex:blind_rotate -s 3 -c 1 -o gpurun_out/${TAG}_${name} $CMD > gpurun_out/${TAG}_${name}_full.log 2>&1
  tail -1 gpurun_out/${TAG}_${name}_full.log
  python tools/ncu_summary.py gpurun_out/${TAG}_${name}.ncu-rep > gpurun_out/${TAG}_${name}.txt 2>&1
  python tools/ncu_stalls.py gpurun_out/${TAG}_${name}.ncu-rep >> gpurun_out/${TAG}_${name}.txt 2>&1
  rm -f gpurun_out/${TAG}_${name}.ncu-rep; head -30 gpurun_out/${TAG}_${name}.txt; }
prof mklwe148 --batch 148
prof mkntru148 --paramset STD100_MKNTRU --batch 148
