#!/bin/bash
# usage: tools/gpu_prof2.sh <tag> -- default bench + ncu (launch list, --set full) of the gate kernel on both k = 2 shapes, and the
# full-batch (4096) capture whose DRAM bytes feed roofline.traffic
TAG=$1; mkdir -p gpurun_out
(time python bench.py --steps 3 --warmup 3) > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
cut -c1-600 gpurun_out/${TAG}_bench.json; tail -2 gpurun_out/${TAG}_bench.err
prof() { name=$1; shift; CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline $*"
  $CMD > gpurun_out/${TAG}_${name}_plain.json 2> gpurun_out/${TAG}_${name}_plain.err || return
  ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/${TAG}_${name}_launches.csv $CMD > /dev/null 2>&1
  $CMD > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:blind_rotate -s 5 -c 1 -o gpurun_out/${TAG}_${name} $CMD > gpurun_out/${TAG}_${name}_full.log 2>&1
  tail -1 gpurun_out/${TAG}_${name}_full.log
  # summarise on the box (the reports of this library exceed the 64 MiB that travel back) and drop the report
  python tools/ncu_summary.py gpurun_out/${TAG}_${name}.ncu-rep > gpurun_out/${TAG}_${name}.txt 2>&1
  python tools/ncu_stalls.py gpurun_out/${TAG}_${name}.ncu-rep >> gpurun_out/${TAG}_${name}.txt 2>&1
  ls -la gpurun_out/${TAG}_${name}.ncu-rep; rm -f gpurun_out/${TAG}_${name}.ncu-rep; }
prof mklwe592 --batch 592
prof mkntru592 --paramset STD100_MKNTRU --batch 592 --synthetic-keys
prof mklwe4096 --batch 4096
