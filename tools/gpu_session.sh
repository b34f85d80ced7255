#!/bin/bash
# usage: tools/gpu_session.sh <tag> [--tests] [--ncu-full] -- one GPU box visit: (tests,) default bench, ncu launch list, (ncu --set full)
TAG=$1; shift
mkdir -p gpurun_out
for a in "$@"; do
  case $a in
    --tests) (time python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_tests.log 2>&1; tail -5 gpurun_out/${TAG}_tests.log;;
  esac
done
(time python bench.py --steps 3 --warmup 3) > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
cut -c1-1500 gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err
CMD="python bench.py --batch 592 --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/${TAG}_plain.json 2> gpurun_out/${TAG}_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_launches.log 2>&1
for a in "$@"; do
  case $a in
    --ncu-full) $CMD > gpurun_out/${TAG}_plain2.json 2> gpurun_out/${TAG}_plain2.err && \
      ncu --set full --clock-control none --import-source on -k regex:blind_rotate -s 5 -c 1 -o gpurun_out/${TAG}_blind_rotate $CMD > gpurun_out/${TAG}_full.log 2>&1
      tail -2 gpurun_out/${TAG}_full.log;;
  esac
done
