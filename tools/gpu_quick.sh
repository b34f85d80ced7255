#!/bin/bash
# usage: tools/gpu_quick.sh <tag> [paramset:batches:check ...] -- quick timings + oracle checks (tests/dev_quick_bench.py)
TAG=$1; shift; mkdir -p gpurun_out; L=gpurun_out/${TAG}_quick.log; : > $L
[ $# -eq 0 ] && set -- STD100_MKNTRU_LWE:592,4096:2 STD100_MKNTRU:592,4096:2 STD100_MKNTRU_2:444,1776:1
for spec in "$@"; do IFS=: read ps b c <<< "$spec"; python tests/dev_quick_bench.py --paramset $ps --batches $b --check $c >> $L 2>&1; done
grep -v "^$" $L | tail -30
