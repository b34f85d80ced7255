#!/bin/bash
# Round-2 GPU sessions, one script: tools/gpu_r2.sh <session> <tag>.  Everything lands in gpurun_out/<tag>_*.
#   prof   ncu --set full of the gate kernels at k = 4 / 8 / 16, the 128-register k = 2 build, the batched transform; per-line
# (transform-loader names: up to session r2s13 "tma" meant one polynomial per ring slot and "2" two; since then "tma" is the
# two-polynomial default and "1" the one-polynomial kernel)
#          shared-memory conflict attribution of the k = 2 gate kernel; integer-pipe peaks
S=$1; TAG=$2; mkdir -p gpurun_out
prof() { # prof <name> <kernel regex> <skip> <count> <bench args...>
  name=$1; rx=$2; skip=$3; cnt=$4; shift 4
  CMD="python bench.py --no-cpu-baseline $*"
  $CMD > gpurun_out/${TAG}_${name}_plain.json 2> gpurun_out/${TAG}_${name}_plain.err || { echo "$name: plain run failed"; tail -3 gpurun_out/${TAG}_${name}_plain.err; return; }
  cut -c1-260 gpurun_out/${TAG}_${name}_plain.json
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c $cnt -o gpurun_out/${TAG}_${name} $CMD > gpurun_out/${TAG}_${name}_full.log 2>&1
  python tools/ncu_summary.py gpurun_out/${TAG}_${name}.ncu-rep > gpurun_out/${TAG}_${name}.txt 2>&1
  python tools/ncu_stalls.py gpurun_out/${TAG}_${name}.ncu-rep >> gpurun_out/${TAG}_${name}.txt 2>&1
  python tools/ncu_lines.py gpurun_out/${TAG}_${name}.ncu-rep 30 >> gpurun_out/${TAG}_${name}.txt 2>&1
  rm -f gpurun_out/${TAG}_${name}.ncu-rep; }
peaks() {
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/pipe_microbench tools/pipe_microbench.cu && /tmp/pipe_microbench > gpurun_out/${TAG}_pipe_microbench.json 2> gpurun_out/${TAG}_pipe_microbench.err
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/bfly_microbench tools/bfly_microbench.cu && /tmp/bfly_microbench > gpurun_out/${TAG}_bfly_microbench.txt 2>&1
  nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks.mem,power.draw --format=csv > gpurun_out/${TAG}_smi.txt; }
run() { name=$1; shift; (time timeout 1500 python bench.py "$@") > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err; \
        echo "== $name"; cut -c1-600 gpurun_out/${TAG}_${name}.json; tail -2 gpurun_out/${TAG}_${name}.err; }
case $S in
prof4)  # ncu of the one-direction transform kernels: every launch of the short run (the summary keeps the full-batch ones)
  prof ntt2048split ntt_batch_tma2 0 8 --workload ntt --ntt-n 2048 --steps 1 --warmup 1 --ntt-variants tma --min-seconds 0
  ;;
prof3)  # ncu of the last kernels of round 2: the one-direction transform kernel, the un-mixed XZW gate kernel
  prof ntt2048fwd ntt_batch_tma2 2 1 --workload ntt --ntt-n 2048 --steps 1 --warmup 1 --ntt-variants tma --min-seconds 0
  prof ntt2048inv ntt_batch_tma2 3 1 --workload ntt --ntt-n 2048 --steps 1 --warmup 1 --ntt-variants tma --min-seconds 0
  prof mkntru4096 blind_rotate_fast 3 1 --steps 1 --warmup 3 --no-also --check-oracle 0 --paramset STD100_MKNTRU
  ;;
final3) # after the un-mixing of the paired step: XZW register budget re-check, party sweep k = 2 / 4 / 8 at batch 16384, then what the driver runs
  q() { echo "== $*"; env "$@" timeout 300 python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 3 2>&1 | grep -E "batch|MISMATCH|exact|rror" ; }
  PS=STD100_MKNTRU BT=4096 CK=1; q X=0; q MKFHE_MINB=4
  run k2  --paramset STD100_MKNTRU   --batch 16384 --steps 2 --warmup 1 --synthetic-keys --no-also --check-oracle 16 --no-cpu-baseline
  run k4  --paramset STD100_MKNTRU_2 --batch 16384 --steps 1 --warmup 1 --synthetic-keys --no-also --check-oracle 16 --no-cpu-baseline
  run k8  --paramset STD100_MKNTRU_3 --batch 16384 --steps 1 --warmup 1 --synthetic-keys --no-also --check-oracle 4 --no-cpu-baseline
  (time python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_tests.log 2>&1; tail -4 gpurun_out/${TAG}_tests.log
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -3 gpurun_out/${TAG}_smoke.log
  run ref --impl reference --steps 3 --warmup 1
  run bench --steps 20 --warmup 3
  ;;
pair2)  # the 80-register XZW build without the dead prefetch code: paired inverses on / off at k = 2 / 4 / 8
  q() { echo "== $*"; env "$@" timeout 300 python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 3 2>&1 | grep -E "batch|MISMATCH|exact|rror" ; }
  PS=STD100_MKNTRU BT=444,4096 CK=1; q MKFHE_PAIR=0; q MKFHE_PAIR=1
  PS=STD100_MKNTRU_2 BT=888 CK=1; q MKFHE_PAIR=0; q MKFHE_PAIR=1
  PS=STD100_MKNTRU_3 BT=592 CK=1; q MKFHE_PAIR=0; q MKFHE_PAIR=1
  PS=STD100_MKNTRU_LWE_2 BT=444 CK=1; q MKFHE_PAIR=0; q MKFHE_PAIR=1
  PS=STD100_MKNTRU_LWE BT=296,444 CK=1; q MKFHE_PAIR=0; q MKFHE_PAIR=1
  ;;
final2) # what the driver runs at round end: GPU tests, smoke, both bench arms at N = 1 with the driver's flags
  (time python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_tests.log 2>&1; tail -4 gpurun_out/${TAG}_tests.log
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -3 gpurun_out/${TAG}_smoke.log
  run ref --impl reference --steps 3 --warmup 1
  run bench --steps 20 --warmup 3
  ;;
cl5)    # cluster kernel with the key words preloaded into registers: latency at k = 2 / 4 / 8 / 16
  q() { echo "== $*"; env "$@" timeout 300 python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 3 2>&1 | grep -E "batch|MISMATCH|exact|rror" ; }
  PS=STD100_MKNTRU_LWE BT=1,74 CK=2; q X=0
  PS=STD100_MKNTRU BT=1,74 CK=2; q X=0
  PS=STD100_MKNTRU_2 BT=1,37 CK=1; q X=0
  PS=STD100_MKNTRU_LWE_2 BT=1 CK=1; q X=0
  PS=STD100_MKNTRU_3 BT=1 CK=1; q X=0
  PS=STD100_MKNTRU_4 BT=1 CK=1; q X=0
  ;;
cl4)    # cluster kernel after the relaxed owner arrival; then the whole GPU suite and smoke with clusters as the small-batch default
  q() { echo "== $*"; env "$@" timeout 300 python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 3 2>&1 | grep -E "batch|MISMATCH|exact|rror" ; }
  PS=STD100_MKNTRU_LWE BT=1,74 CK=1; q X=0
  PS=STD100_MKNTRU BT=1,74 CK=1; q X=0
  PS=STD100_MKNTRU_2 BT=1,37 CK=1; q X=0
  (time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_tests.log 2>&1; tail -5 gpurun_out/${TAG}_tests.log
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -3 gpurun_out/${TAG}_smoke.log
  ;;
cl3)    # cluster kernel: L1 key prefetch on / off (MKFHE_PREFETCH), ncu of the kernel, default bench line with the single-gate block
  q() { echo "== $*"; env "$@" timeout 300 python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 3 2>&1 | grep -E "batch|MISMATCH|exact|rror" ; }
  PS=STD100_MKNTRU_LWE BT=1,74 CK=1; q MKFHE_PREFETCH=0; q X=0
  PS=STD100_MKNTRU BT=1,74 CK=1; q MKFHE_PREFETCH=0; q X=0
  PS=STD100_MKNTRU_2 BT=1 CK=1; q MKFHE_PREFETCH=0; q X=0
  prof cluster74 blind_rotate_cluster 2 1 --batch 74 --steps 1 --warmup 2 --synthetic-keys --no-also --check-oracle 0
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -3 gpurun_out/${TAG}_smoke.log
  run bench --steps 5 --warmup 3
  ;;
cl2)    # cluster sizes 2 / 4 / 8 at k = 4 / 8 / 16 (single gate and the largest batch that still fits), then the kernel-variant tests
  q() { echo "== $*"; env "$@" timeout 300 python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 3 2>&1 | grep -E "batch|MISMATCH|exact|rror" ; }
  PS=STD100_MKNTRU_2 BT=1,16 CK=1; q MKFHE_CLUSTER=2; q MKFHE_CLUSTER=4; q X=0
  PS=STD100_MKNTRU_LWE_2 BT=1,16 CK=1; q MKFHE_CLUSTER=4
  PS=STD100_MKNTRU_3 BT=1,16 CK=1; q MKFHE_CLUSTER=2; q MKFHE_CLUSTER=4; q MKFHE_CLUSTER=8; q X=0
  PS=STD100_MKNTRU_4 BT=1,16 CK=1; q MKFHE_CLUSTER=2; q MKFHE_CLUSTER=4; q MKFHE_CLUSTER=8; q X=0
  PS=STD100_MKNTRU_LWE BT=1,37,74,75,148,149 CK=1; q X=0
  (time timeout 900 python -m pytest tests/test_gpu_partysweep.py -m gpu -x -q) > gpurun_out/${TAG}_partysweep.log 2>&1; tail -5 gpurun_out/${TAG}_partysweep.log
  ;;
cl)     # two SMs per gate (blind_rotate_cluster) against the party-parallel kernel on one SM, small batches
  q() { echo "== $*"; env "$@" timeout 300 python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 3 2>&1 | grep -E "batch|MISMATCH|exact|Error|error|rror" ; }
  PS=STD100_MKNTRU_LWE BT=1,16,74 CK=2; q MKFHE_CLUSTER=0; q MKFHE_CLUSTER=1
  PS=STD100_MKNTRU BT=1,16,74 CK=2; q MKFHE_CLUSTER=0; q MKFHE_CLUSTER=1
  PS=STD100_MKNTRU_2 BT=1,74 CK=2; q MKFHE_CLUSTER=0; q MKFHE_CLUSTER=1
  PS=STD100_MKNTRU_LWE_2 BT=1,74 CK=2; q MKFHE_CLUSTER=0; q MKFHE_CLUSTER=1
  PS=STD100_MKNTRU_3 BT=74 CK=1; q MKFHE_CLUSTER=0; q MKFHE_CLUSTER=1
  PS=STD100_MKNTRU_4 BT=74 CK=1; q MKFHE_CLUSTER=0; q MKFHE_CLUSTER=1
  ;;
ag)     # accumulators in global memory (blind_rotate_fast_ag) against the shared-memory kernels, k = 4 / 8 / 16, both flavours
  q() { echo "== $*"; env "$@" python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 2 2>&1 | grep -E "batch|MISMATCH|exact|Error|error" ; }
  PS=STD100_MKNTRU_4 BT=592,1184 CK=1; q MKFHE_AG=0; q MKFHE_AG=1; q MKFHE_AG=1 MKFHE_MINB=4
  PS=STD100_MKNTRU_LWE_4 BT=592,1184 CK=1; q MKFHE_AG=0; q MKFHE_AG=1; q MKFHE_AG=1 MKFHE_MINB=3
  PS=STD100_MKNTRU_3 BT=592,1184 CK=1; q MKFHE_AG=0; q MKFHE_AG=1; q MKFHE_AG=1 MKFHE_MINB=4
  PS=STD100_MKNTRU_LWE_3 BT=592,1184 CK=1; q MKFHE_AG=0; q MKFHE_AG=1
  PS=STD100_MKNTRU_2 BT=888,1776 CK=1; q MKFHE_AG=0; q MKFHE_AG=1
  PS=STD100_MKNTRU_LWE_2 BT=888,1776 CK=1; q MKFHE_AG=0; q MKFHE_AG=1
  PS=STD100_MKNTRU_LWE BT=1184 CK=1; q MKFHE_AG=0; q MKFHE_AG=1
  ;;
noalloc)  # the allocation-free circuit entry point: its test alone in a fresh process, then after the whole suite, then smoke
  (time python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q) > gpurun_out/${TAG}_noalloc_alone.log 2>&1; tail -4 gpurun_out/${TAG}_noalloc_alone.log
  (time python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_tests.log 2>&1; tail -4 gpurun_out/${TAG}_tests.log
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -3 gpurun_out/${TAG}_smoke.log
  ;;
check)  # full GPU test suite, smoke, default bench line, both transform loaders at N = 2048 / 1024
  (time python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_tests.log 2>&1; tail -5 gpurun_out/${TAG}_tests.log
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -2 gpurun_out/${TAG}_smoke.log
  run bench --steps 5 --warmup 3
  run ntt2048 --workload ntt --ntt-n 2048 --steps 20 --warmup 3 --ntt-variants tma,1,ldg
  run ntt1024 --workload ntt --ntt-n 1024 --steps 20 --warmup 3 --ntt-variants tma,1,ldg
  ;;
prof2)  # final kernels: launch list of the default bench command, ncu --set full of the gate kernels on the bench's own launches
        # (distinct gates, reference keys), the TMA transform kernel, the party-parallel kernel with the held-product schedule
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --check-oracle 0"
  $CMD > gpurun_out/${TAG}_bench_plain.json 2> gpurun_out/${TAG}_bench_plain.err && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_bench_launches.csv $CMD > gpurun_out/${TAG}_bench_launches.log 2>&1
  prof mklwe4096  blind_rotate_fast 3 1 --steps 1 --warmup 3 --no-also --check-oracle 0
  prof mkntru4096 blind_rotate_fast 3 1 --steps 1 --warmup 3 --no-also --check-oracle 0 --paramset STD100_MKNTRU
  prof ntt2048tma ntt_batch_tma 4 2 --workload ntt --ntt-n 2048 --steps 1 --warmup 1 --ntt-variants tma --min-seconds 0
  prof mklwe148   blind_rotate_grouped 3 1 --batch 148 --steps 1 --warmup 3 --synthetic-keys --no-also --check-oracle 0
  prof k16_148    blind_rotate_grouped 1 1 --paramset STD100_MKNTRU_4 --batch 148 --steps 1 --warmup 1 --synthetic-keys --check-oracle 0
  # calibration: what the pipe counters read while the GPU does nothing but register-resident butterflies (the roofline's peak)
  prof bflypeak   butterfly_peak 1 1 --batch 148 --steps 1 --warmup 1 --synthetic-keys --no-also --check-oracle 0
  ;;
ntt2)   # full GPU tests + smoke, then the transform loaders incl. the two-polynomial ring
  (time python -m pytest tests -m gpu -q) > gpurun_out/${TAG}_tests.log 2>&1; tail -4 gpurun_out/${TAG}_tests.log
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -3 gpurun_out/${TAG}_smoke.log
  run ntt2048 --workload ntt --ntt-n 2048 --steps 20 --warmup 3 --ntt-variants tma,2,ldg
  run ntt1024 --workload ntt --ntt-n 1024 --steps 20 --warmup 3 --ntt-variants tma,2,ldg
  ;;
final)  # what the driver runs at round end (tests, smoke, both bench arms at N = 1), after one more A/B
  q() { echo "== $*"; env "$@" python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 3 2>&1 | grep -E "batch|MISMATCH|exact|Error|error" ; }
  PS=STD100_MKNTRU_LWE BT=592,4096 CK=1; q X=0; q MKFHE_FAST_ALIAS=1
  (time python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_tests.log 2>&1; tail -4 gpurun_out/${TAG}_tests.log
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -3 gpurun_out/${TAG}_smoke.log
  run ref --impl reference --steps 3 --warmup 1
  run bench --steps 20 --warmup 3
  ;;
sweep)  # configs[2]: party-count sweep at batch 16384 (synthetic keys), oracle sample 16/16/4/1, CPU leg = the C port on all cores
  q() { echo "== $*"; env "$@" python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 3 2>&1 | grep -E "batch|MISMATCH|exact|Error|error" ; }
  PS=STD100_MKNTRU_LWE BT=4096 CK=1; q X=0; q MKFHE_MINB=3
  PS=STD100_MKNTRU_LWE_2 BT=888,4096 CK=1; q X=0
  run k2  --paramset STD100_MKNTRU   --batch 16384 --steps 2 --warmup 1 --synthetic-keys --no-also --check-oracle 16 --cpu-baseline-gates-per-thread 4
  run k4  --paramset STD100_MKNTRU_2 --batch 16384 --steps 1 --warmup 1 --synthetic-keys --no-also --check-oracle 16 --cpu-baseline-gates-per-thread 1
  run k8  --paramset STD100_MKNTRU_3 --batch 16384 --steps 1 --warmup 1 --synthetic-keys --no-also --check-oracle 4 --cpu-baseline-gates-per-thread 1
  run k16 --paramset STD100_MKNTRU_4 --batch 16384 --steps 1 --warmup 1 --synthetic-keys --no-also --check-oracle 1 --cpu-baseline-gates-per-thread 1
  ;;
pair)   # paired inverse transforms (MKFHE_PAIR=0/1) in the 80- / 128-register builds of the sequential kernel
  q() { echo "== $*"; env "$@" python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 3 2>&1 | grep -E "batch|MISMATCH|exact|Error|error" ; }
  PS=STD100_MKNTRU_3 BT=296,592 CK=2; q MKFHE_PAIR=0; q MKFHE_PAIR=1
  PS=STD100_MKNTRU_2 BT=444,888 CK=2; q MKFHE_PAIR=0; q MKFHE_PAIR=1
  PS=STD100_MKNTRU_LWE_2 BT=444,888 CK=2; q MKFHE_PAIR=0; q MKFHE_PAIR=1; q MKFHE_PAIR=1 MKFHE_FAST_ALIAS=1; q MKFHE_PAIR=0 MKFHE_FAST_ALIAS=1
  PS=STD100_MKNTRU_LWE BT=296,444 CK=2; q MKFHE_PAIR=0; q MKFHE_PAIR=1
  PS=STD100_MKNTRU BT=296,444,4096 CK=2; q MKFHE_PAIR=0; q MKFHE_PAIR=1
  ;;
exp2)   # party-parallel groups after the held-product schedule: 1 / 2 / 3 groups at k = 8, 16 (both flavours), k = 4
  q() { echo "== $*"; env "$@" python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 3 2>&1 | grep -E "batch|MISMATCH|exact|Error|error" ; }
  PS=STD100_MKNTRU_3 BT=148,296,592 CK=1; q X=0; q MKFHE_GROUPS=2; q MKFHE_GROUPS=3
  PS=STD100_MKNTRU_LWE_3 BT=148,296,592 CK=1; q X=0; q MKFHE_GROUPS=2; q MKFHE_GROUPS=3
  PS=STD100_MKNTRU_4 BT=148,296 CK=1; q X=0; q MKFHE_GROUPS=3
  PS=STD100_MKNTRU_2 BT=148,444 CK=1; q X=0; q MKFHE_GROUPS=3
  PS=STD100_MKNTRU_LWE_2 BT=148,444 CK=1; q X=0; q MKFHE_GROUPS=3
  ;;
lat)    # small batches: L2 key prefetch on / off (MKFHE_PREFETCH), then the config-4 job on one GPU
  (time python -m pytest tests -m gpu -x -q) > gpurun_out/${TAG}_tests.log 2>&1; tail -3 gpurun_out/${TAG}_tests.log
  q() { echo "== $*"; env "$@" python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 5 2>&1 | grep -E "batch|MISMATCH|exact|Error|error" ; }
  PS=STD100_MKNTRU_LWE BT=1,16,148,296 CK=1; q MKFHE_PREFETCH=0; q MKFHE_PREFETCH=1
  PS=STD100_MKNTRU BT=1,148,296 CK=1; q MKFHE_PREFETCH=0; q MKFHE_PREFETCH=1
  PS=STD100_MKNTRU_2 BT=1,148 CK=1; q MKFHE_PREFETCH=0; q MKFHE_PREFETCH=1
  PS=STD100_MKNTRU_4 BT=148 CK=1; q MKFHE_PREFETCH=0; q MKFHE_PREFETCH=1
  bash tools/gpu_r2.sh sharded $TAG 1 1048576
  ;;
nccl)   # transports and collective bandwidth of this box: tools/gpu_r2.sh nccl <tag> <N>
  N=${3:-2}
  NCCL_DEBUG=INFO python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tools/nccl_probe.py > gpurun_out/${TAG}_nccl_n$N.log 2>&1
  grep -E "^\{" gpurun_out/${TAG}_nccl_n$N.log > gpurun_out/${TAG}_nccl_n$N.json
  grep -E "via |NVLS|P2P|SHM|Channel 00|Connected|nvlink|NVLink" gpurun_out/${TAG}_nccl_n$N.log | head -40 > gpurun_out/${TAG}_nccl_n${N}_transports.txt
  nvidia-smi topo -m > gpurun_out/${TAG}_topo.txt 2>&1
  cat gpurun_out/${TAG}_nccl_n$N.json; head -12 gpurun_out/${TAG}_nccl_n${N}_transports.txt; cat gpurun_out/${TAG}_topo.txt | head -14
  ;;
multi)  # N GPUs (gpurun --gpus N): default bench under torchrun, both arms; then the config-4 job at --total-gates $3 (default 2^17)
  N=${3:-2}; TG=${4:-131072}
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
  (time timeout 1200 $TR bench.py --gpus $N --steps 5 --warmup 3) > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err; cut -c1-300 gpurun_out/${TAG}_bench_n$N.json; tail -3 gpurun_out/${TAG}_bench_n$N.err
  (time timeout 1500 $TR bench.py --gpus $N --workload sharded --paramset STD100_MKNTRU_2 --total-gates $TG --steps 1 --warmup 1 --check-oracle 16) > gpurun_out/${TAG}_sharded_n$N.json 2> gpurun_out/${TAG}_sharded_n$N.err; cut -c1-300 gpurun_out/${TAG}_sharded_n$N.json; tail -3 gpurun_out/${TAG}_sharded_n$N.err
  ;;
n8)     # 8 GPUs: the config-4 job, then both arms of the default bench under torchrun (what the driver runs at round end)
  bash tools/gpu_r2.sh sharded $TAG 8 1048576
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29514"
  (time timeout 900 $TR bench.py --impl reference --gpus 8 --steps 2 --warmup 1) > gpurun_out/${TAG}_ref_n8.json 2> gpurun_out/${TAG}_ref_n8.err; cut -c1-300 gpurun_out/${TAG}_ref_n8.json
  (time timeout 900 $TR bench.py --gpus 8 --steps 5 --warmup 3) > gpurun_out/${TAG}_bench_n8.json 2> gpurun_out/${TAG}_bench_n8.err; cut -c1-300 gpurun_out/${TAG}_bench_n8.json; tail -3 gpurun_out/${TAG}_bench_n8.err
  ;;
sharded)  # the config-4 job alone: tools/gpu_r2.sh sharded <tag> <N> <total gates>
  N=${3:-1}; TG=${4:-1048576}
  if [ "$N" = 1 ]; then TR="python"; else TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"; fi
  (time timeout 1700 $TR bench.py --gpus $N --workload sharded --paramset STD100_MKNTRU_2 --total-gates $TG --steps 1 --warmup 1 --check-oracle 16) > gpurun_out/${TAG}_sharded_n$N.json 2> gpurun_out/${TAG}_sharded_n$N.err; cut -c1-400 gpurun_out/${TAG}_sharded_n$N.json; tail -3 gpurun_out/${TAG}_sharded_n$N.err
  ;;
exp1)   # experiments: transform-kernel occupancy, CTA de-phasing, shuffle vs shared-memory exchange, 3 party-parallel groups
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/exchange_microbench tools/exchange_microbench.cu && /tmp/exchange_microbench > gpurun_out/${TAG}_exchange_microbench.txt 2>&1
  cat gpurun_out/${TAG}_exchange_microbench.txt
  for c in 4 5 6; do MKFHE_NTT_CTAS=$c run ntt2048_ctas$c --workload ntt --ntt-n 2048 --steps 20 --warmup 3 --ntt-variants tma --min-seconds 1.5; done
  for c in 4 5 6; do MKFHE_NTT_CTAS=$c run ntt1024_ctas$c --workload ntt --ntt-n 1024 --steps 20 --warmup 3 --ntt-variants tma --min-seconds 1.5; done
  q() { echo "== $*"; env "$@" python tests/dev_quick_bench.py --paramset $PS --batches $BT --check $CK --reps 3 2>&1 | grep -E "batch|MISMATCH|exact|Error|error" ; }
  PS=STD100_MKNTRU_LWE BT=4096 CK=2
  q X=0; q MKFHE_STAGGER_NS=3000; q MKFHE_STAGGER_NS=6000; q MKFHE_STAGGER_NS=12000
  PS=STD100_MKNTRU_3 BT=148,296,592 CK=1
  q X=0; q MKFHE_GROUPS=3; q MKFHE_GROUPS=2
  PS=STD100_MKNTRU_4 BT=148,296 CK=1
  q X=0; q MKFHE_GROUPS=3; q MKFHE_GROUPS=1
  PS=STD100_MKNTRU_2 BT=148,444,888 CK=1
  q X=0; q MKFHE_GROUPS=3
  ;;
prof)
  peaks
  prof k4_444   blind_rotate 1 1 --paramset STD100_MKNTRU_2 --batch 444 --steps 1 --warmup 1 --synthetic-keys
  prof k8_296   blind_rotate 1 1 --paramset STD100_MKNTRU_3 --batch 296 --steps 1 --warmup 1 --synthetic-keys
  prof k16_148  blind_rotate 1 1 --paramset STD100_MKNTRU_4 --batch 148 --steps 1 --warmup 1 --synthetic-keys
  prof mklwe296 blind_rotate 3 1 --batch 296 --steps 1 --warmup 3 --synthetic-keys
  prof mklwe1184 blind_rotate 3 1 --batch 1184 --steps 1 --warmup 3 --synthetic-keys
  prof ntt2048  ntt_batch 2 2 --workload ntt --ntt-n 2048 --steps 1 --warmup 1
  prof ntt1024  ntt_batch 2 2 --workload ntt --ntt-n 1024 --steps 1 --warmup 1
  ;;
esac
ls -la gpurun_out | head -50
