#!/usr/bin/env python
"""bench.py -- multi-key NAND bootstraps/sec on B200 (the metric of BASELINE.json), beside the reference's CPU path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--paramset NAME] [--batch B]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A "step" is one pass of the hot path (BinFHEContext::EvalBinGate(NAND) = input combine + blind rotation + extraction +
modulus switch + key switch) over one batch of gates.  Default workload = BASELINE.json configs[1]: boolean-mklwe shape
(STD100_MKNTRU_LWE, k = 2 parties), batch 4096 per GPU.  Multi-GPU: each rank holds a replica of all keys (rank 0
generates, NCCL broadcast) and bootstraps its own batch (weak scaling, no data-path collective); results are gathered
after the timed region.

One JSON line on stdout (rank 0).  `value` = gates/s with inputs resident in HBM; `e2e` = the same through the C ABI with
host buffers (H2D + D2H inside the timed region; N > 1: + the result gather); `roofline` = modular multiplications per second
of the blind-rotation kernel (SURVEY.md section 8d count) against the lazy-butterfly rate measured on this GPU in this run,
and against the analytic 8-cycle floor; `roofline_hbm` = key-stream bytes against MEASURED_PEAKS.json; `cpu_baseline` = the
reference's own code (oracle/_ref, compiled unmodified) on all host cores over a bounded sample, plus the single-thread
single-gate latency the reference's examples print; `single_gate` = the same one-gate call on the GPU (BASELINE.json configs[0]:
kernel ms, host-buffer call ms, SMs per gate of the thread-block cluster it runs on).  The default workload adds `also.boolean_mkntru`: the same measurement
for STD100_MKNTRU (the XZW accumulator of mk-acc-xzw.cpp), k = 2, same batch.  Parity inside the run (outside the timed
regions): every output decrypted = NAND, 64 gates bit-exact against the CPU oracle.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "multi-key NAND bootstraps/sec"
UNIT = "gates/s"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


_REAL_STDOUT = None


def capture_stdout():
    """Libraries (NCCL's version banner, for one) write to fd 1; the contract is ONE JSON line on stdout.  Point fd 1 at
    stderr for the duration of the run and keep the real stdout for emit()."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


# ---------------------------------------------------------------- algorithmic work
def work_model(P):
    """modmul/gate and cost-model instructions/gate: S*(T*beta + Pw) with 7 instr per butterfly, 5 per pointwise MAC."""
    k, N, n, ell = P.k, P.N, P.n, P.digitsG - 1
    S = k * n
    T = (k + 1) * (ell + 1)                 # transforms per step (= 3k+3 at ell = 2)
    beta = (N // 2) * (N.bit_length() - 1)  # butterflies per transform
    if P.variant == 0:                      # XZW: k monomial + 2*ell key assembly + (2k+1)*ell HbProd
        Pw = (k + 2 * ell + (2 * k + 1) * ell) * N
        key_bytes_step = 2 * ell * 2 * N * 4
    else:                                   # XZW_B
        Pw = (k + (2 * k + 1) * ell) * N
        key_bytes_step = ell * 2 * N * 4
    return dict(steps=S, transforms=S * T, modmul=S * (T * beta + Pw), instr=S * (T * beta * 7 + Pw * 5),
                key_bytes=S * key_bytes_step)


def ncu_traffic(paramset, batch):
    """DRAM bytes (read + write) of one gate-kernel launch, from the committed ncu --set full capture of the same workload
    (profiles/ncu_traffic.json, written from profiles/r1_*_<workload>.txt); None when no capture matches this launch shape."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        return t.get(f"{paramset}:{batch}", {}).get("dram_bytes")
    except Exception:
        return None


# ---------------------------------------------------------------- clocks sampling (B200_PROFILING.md recipe)
class ClockSampler:
    """nvidia-smi clocks / throttle reasons every 100 ms while a timed region runs (the recipe's clocks line), plus one
    synchronous sample at start() and one at stop(), so a region shorter than the sampling period still has a record; arms
    whose K steps would take well under a second repeat them until the region is long enough to be sampled under load."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc, self.edge = index, [], None, []

    def _once(self):
        try:
            out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                 capture_output=True, text=True, timeout=20).stdout
            for line in out.splitlines():
                self.edge.append([x.strip() for x in line.split(",")])
        except Exception as e:
            log("clock sampler: one-shot query failed:", e)

    def start(self):
        self._once()
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception as e:  # nvidia-smi missing: report it, do not fake numbers
            log("clock sampler unavailable:", e)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        self._once()
        rows = self.rows if self.rows else self.edge
        sm = sorted(int(r[1]) for r in rows if len(r) > 2 and r[1].isdigit())
        mx = [int(r[2]) for r in rows + self.edge if len(r) > 2 and r[2].isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows + self.edge:
            for nme, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx[0] if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "samples_during": len(self.rows), "samples_edge": len(self.edge)}


# ---------------------------------------------------------------- workload (keys + ciphertexts)
def make_workload(paramset: str, method: str, batch: int, seed: int, encrypt: bool = True):
    """Keys and input ciphertexts.  With the prebuilt reference present (oracle/_ref) they are the reference's own keygen and
    Encrypt output, so a sample of results can be decrypted and checked = NAND after the run; otherwise random residues of
    the right shapes (throughput is data-independent).  Returns (Params, keys dict, ct1, b1, ct2, b2, bits or None, ref)."""
    import mkfhe_b200 as mk
    from oracle import refapi

    P = mk.derive_params(mk.PARAMSETS[paramset], method)
    rng = np.random.default_rng(seed)
    lwe = P.scheme == 1
    if refapi.available() and paramset in refapi.PARAMSETS and not os.environ.get("MKFHE_BENCH_SYNTHETIC_KEYS"):
        t0 = time.time()
        r = refapi.Ref.from_paramset(paramset)
        r.keygen()
        keys = dict(bk=r.bk(), pkey=r.pkey())
        if lwe:
            keys["kskA"], keys["kskB"] = r.ksk_mklwe()
        else:
            keys["ksk"], _ = r.ksk_mntru()
            keys["ctnand"] = r.ctnand()
        if not encrypt:  # the caller builds its own inputs from this context (distinct_gate_inputs)
            log(f"[bench] reference keygen: {time.time() - t0:.1f}s (data: reference keys)")
            return P, keys, None, None, None, None, None, r
        nuniq = min(batch, 64)  # distinct fresh encryptions, tiled to the batch
        m0, m1 = rng.integers(0, 2, nuniq), rng.integers(0, 2, nuniq)
        e0 = [r.encrypt(int(m)) for m in m0]
        e1 = [r.encrypt(int(m)) for m in m1]
        ct1 = np.array([x[0] for x in e0], dtype=np.uint32).reshape(nuniq, P.k, P.n)
        ct2 = np.array([x[0] for x in e1], dtype=np.uint32).reshape(nuniq, P.k, P.n)
        b1 = np.array([x[1] for x in e0], dtype=np.uint32).reshape(nuniq)
        b2 = np.array([x[1] for x in e1], dtype=np.uint32).reshape(nuniq)
        bits = (m0.astype(np.int64), m1.astype(np.int64))
        log(f"[bench] reference keygen + {2 * nuniq} encryptions: {time.time() - t0:.1f}s (data: reference keys)")
        return P, keys, ct1, b1, ct2, b2, bits, r
    ell, S = P.digitsG - 1, (2 if P.variant == 0 else 1)
    keys = dict(bk=rng.integers(0, P.Q, size=(P.k * S * P.n + 1) * ell * 2 * P.N, dtype=np.uint32),
                pkey=rng.integers(0, P.Q, size=P.k * ell * P.N, dtype=np.uint32))
    if lwe:
        rows = P.k * P.N * P.baseKS * P.dKS
        keys["kskA"] = rng.integers(0, P.qKS, size=rows * P.n, dtype=np.uint32)
        keys["kskB"] = rng.integers(0, P.qKS, size=rows, dtype=np.uint32)
    else:
        keys["ksk"] = rng.integers(0, P.qKS, size=P.k * P.N * P.dKS * P.n, dtype=np.uint32)
        keys["ctnand"] = rng.integers(0, P.q, size=P.k * P.n, dtype=np.uint32)
    ct1 = rng.integers(0, P.q, size=(batch, P.k, P.n), dtype=np.uint32)
    ct2 = rng.integers(0, P.q, size=(batch, P.k, P.n), dtype=np.uint32)
    b1 = rng.integers(0, P.q, size=batch, dtype=np.uint32)
    b2 = rng.integers(0, P.q, size=batch, dtype=np.uint32)
    return P, keys, ct1, b1, ct2, b2, None, None


# ---------------------------------------------------------------- CPU arms
def cpu_gates_per_s(paramset: str, threads: int, gates: int, ref=None, force_port: bool = False):
    """The reference's EvalBinGate(NAND) on `threads` host threads over `gates` gates.  kind = "reference" when
    oracle/_ref (the unmodified reference, compiled in place) is present, else "port" (oracle/mkfhe_oracle.c)."""
    from oracle import refapi

    if not force_port and (ref is not None or (refapi.available() and paramset in refapi.PARAMSETS)):
        r = ref
        if r is None:
            r = refapi.Ref.from_paramset(paramset)
            r.keygen()
        secs = r.bench_gates(gates, threads)
        return gates / secs, "reference", secs
    # port: random keys of the right shape (run time of the oracle does not depend on key values)
    import mkfhe_b200 as mk
    from oracle import portapi

    method = mk.MKNTRU_LWE if "LWE" in paramset else mk.MKNTRU
    P, keys, ct1, b1, ct2, b2, _, _ = make_workload(paramset, method, 1, 99)
    o = portapi.Oracle(k=P.k, N=P.N, Q=P.Q, psi=P.psi, q=P.q, qKS=P.qKS, n=P.n, baseG_log2=P.baseG_log2, digitsG=P.digitsG,
                       baseKS=P.baseKS, dKS=P.dKS, variant=P.variant, scheme=P.scheme)
    secs = o.bench_gates(keys["bk"], keys["pkey"], keys.get("kskA", keys.get("ksk")), keys.get("kskB"), keys.get("ctnand"),
                         ct1[0].ravel(), int(b1[0]), ct2[0].ravel(), int(b2[0]), gates, threads)
    return gates / secs, "port", secs


def reference_gates(paramset, steps, warmup, gates_per_thread):
    """The reference's EvalBinGate(NAND) for `paramset` on all host threads: K steps of threads * gates_per_thread gates."""
    from oracle import refapi

    cores = os.cpu_count() or 1
    threads = cores
    gates_per_step = max(threads, 1) * gates_per_thread
    ref = None
    if refapi.available() and paramset in refapi.PARAMSETS:
        ref = refapi.Ref.from_paramset(paramset)
        ref.keygen()
    for _ in range(warmup):
        cpu_gates_per_s(paramset, threads, threads, ref)
    times, kind = [], "port"
    for _ in range(steps):
        gps, kind, secs = cpu_gates_per_s(paramset, threads, gates_per_step, ref)
        times.append(secs)
    total = sum(times)
    value = gates_per_step * steps / total
    sample = f"{steps} steps x {gates_per_step} EvalBinGate(NAND) on {threads} threads, shared keys"
    cb = {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample}
    if ref is not None:
        ms = sorted(ref.time_gate(5))
        cb["single_thread_ms"] = ms[len(ms) // 2]
        cb["single_thread_sample"] = "median of 5 single EvalBinGate(NAND) calls, steady_clock, 1 thread"
    return value, 1e3 * total / steps, kind, cb


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # rank 0 alone runs the CPU arm; the others exit 0 without work
    value, ms_step, kind, cb = reference_gates(args.paramset, args.steps, args.warmup, args.ref_gates_per_thread)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "reference keygen + Encrypt" if kind == "reference" else "synthetic",
            "config": workload_config(args), "cpu_baseline": cb,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if args.paramset == "STD100_MKNTRU_LWE" and not args.no_also:  # the boolean-mkntru block of the GPU arm, on the CPU
        v2, ms2, kind2, cb2 = reference_gates("STD100_MKNTRU", min(args.steps, 2), min(args.warmup, 1), max(1, args.ref_gates_per_thread // 2))
        line["also"] = {"boolean_mkntru": {"impl": "reference", "metric": METRIC, "value": v2, "unit": UNIT, "ms_per_step": ms2,
                                           "config": gates_config("STD100_MKNTRU", args.batch, args.gpus), "cpu_baseline": cb2}}
    emit(line)


def gates_config(paramset, batch, world, distinct=None):
    lwe = "LWE" in paramset
    k = {"": 2, "2": 4, "3": 8, "4": 16}.get(paramset.rsplit("_", 1)[-1] if paramset[-1].isdigit() else "", 2)
    cfg = {"workload": f"{paramset} ({'boolean-mklwe' if lwe else 'boolean-mkntru'} shape, k={k}), NAND gate batch {batch} per GPU",
           "paramset": paramset, "batch_per_gpu": batch, "l2": "L2 flushed (256 MiB write) between timed steps",
           "parallelism": f"gates sharded x{world}, keys replicated"}
    if distinct is not None:
        cfg["distinct_gates"] = distinct
    return cfg


def workload_config(args, batch=None):
    return gates_config(args.paramset, batch if batch is not None else args.batch, args.gpus)


def make_oracle(P):
    from oracle import portapi

    return portapi.Oracle(k=P.k, N=P.N, Q=P.Q, psi=P.psi, q=P.q, qKS=P.qKS, n=P.n, baseG_log2=P.baseG_log2, digitsG=P.digitsG,
                          baseKS=P.baseKS, dKS=P.dKS, variant=P.variant, scheme=P.scheme)


# ---------------------------------------------------------------- GPU arm
def distinct_gate_inputs(ctx, P, ref, rng, B):
    """B DISTINCT gates from 2 * nu reference encryptions (nu = min(64, ceil(sqrt(B)))): all nu x nu pairings.
    boolean-mklwe: the paired ciphertexts are REFRESHED ones (outputs of a first NAND level computed here on the GPU) -- the
    inputs every interior gate of a circuit sees, dense mod q; the reference's MKLWE Encrypt samples `a` from a narrow
    Gaussian (mklwe-pke.cpp:53-54), so fresh ciphertexts rotate by ~0 everywhere, a degenerate input.  boolean-mkntru: fresh
    encryptions are dense already (ct = (m q/4 + e) F^-1, mntru-pke.cpp:158-205) and are paired directly; refreshed MKNTRU
    outputs do not compose in the reference itself, so a second level would not decrypt to NAND.
    Returns ct1, b1, ct2, b2 [B...] and the expected plaintext of every gate."""
    lwe = P.scheme == 1
    nu = min(64, max(2, int(np.ceil(np.sqrt(B)))))
    m0, m1 = rng.integers(0, 2, nu).astype(np.int64), rng.integers(0, 2, nu).astype(np.int64)
    e0 = [ref.encrypt(int(m)) for m in m0]
    e1 = [ref.encrypt(int(m)) for m in m1]
    f1 = np.array([x[0] for x in e0], dtype=np.uint32).reshape(nu, P.k, P.n)
    f2 = np.array([x[0] for x in e1], dtype=np.uint32).reshape(nu, P.k, P.n)
    fb1 = np.array([x[1] for x in e0], dtype=np.uint32)
    fb2 = np.array([x[1] for x in e1], dtype=np.uint32)
    if lwe:
        r1 = lambda a: np.roll(a, 1, axis=0)
        x, bx = ctx.EvalBinGate("NAND", f1, f2, fb1, fb2)
        y, by = ctx.EvalBinGate("NAND", r1(f1), f2, r1(fb1), fb2)
        mx, my = 1 - (m0 & m1), 1 - (r1(m0) & m1)
    else:
        x, bx, y, by, mx, my = f1, fb1, f2, fb2, m0, m1
    g = np.arange(B) % (nu * nu)
    i, j = g // nu, g % nu
    c = np.ascontiguousarray
    return c(x[i]), c(np.asarray(bx)[i]), c(y[j]), c(np.asarray(by)[j]), (1 - (mx[i] & my[j])).astype(np.int64), nu * nu


def check_against_oracle(P, keys, ct1, b1, ct2, b2, out, bout, count):
    """Bit-exact check of `count` gates (first, last, evenly spread) against the CPU oracle on all host threads (the C
    library releases the GIL), outside the timed region.  The oracle is the checker here, never the thing measured."""
    if not count:
        return 0
    from concurrent.futures import ThreadPoolExecutor

    o = make_oracle(P)
    B = len(out)
    idx = sorted(set(int(x) for x in np.linspace(0, B - 1, count)))
    t0 = time.time()

    def one(i):
        eo, eb = o.eval_nand(keys["bk"], keys["pkey"], keys.get("kskA", keys.get("ksk")), keys.get("kskB"), keys.get("ctnand"),
                             np.ascontiguousarray(ct1[i]).ravel(), int(b1[i]) if b1 is not None else 0,
                             np.ascontiguousarray(ct2[i]).ravel(), int(b2[i]) if b2 is not None else 0)
        return bool((out[i].ravel() == eo).all() and (bout is None or int(bout[i]) == eb))

    with ThreadPoolExecutor(max_workers=os.cpu_count() or 1) as ex:
        ok = list(ex.map(one, idx))
    if not all(ok):
        raise SystemExit(f"bench.py: gate {idx[ok.index(False)]} differs from the oracle")
    log(f"[bench] {len(idx)} gates bit-exact vs the CPU oracle ({time.time() - t0:.1f}s)")
    return len(idx)


class GpuEnv:
    """Process-wide state of the GPU arm: rank, device, communicator, one side stream, the L2 flush buffer."""

    def __init__(self):
        import torch

        self.torch = torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the GPU arm has no CPU fallback (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.dist, self.nccl_init_s = None, 0.0
        if self.world > 1:
            import torch.distributed as dist

            t0 = time.time()
            dist.init_process_group("nccl", device_id=self.dev)
            w = torch.ones(1 << 20, device=self.dev)
            dist.all_reduce(w)  # the first collective builds the communicator (rings / NVLS trees) ...
            dist.broadcast(w, 0)  # ... and the first broadcast / gather their own connections and buffers
            dist.gather(w, [torch.empty_like(w) for _ in range(self.world)] if self.rank == 0 else None, dst=0)
            torch.cuda.synchronize()
            self.nccl_init_s = time.time() - t0
            self.dist = dist
        self.stream = torch.cuda.Stream(device=self.dev)  # an explicit stream for torch and the library alike
        torch.cuda.set_stream(self.stream)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)
        self.smi_index = self.local if os.environ.get("CUDA_VISIBLE_DEVICES") is None else \
            int(os.environ["CUDA_VISIBLE_DEVICES"].split(",")[self.local])

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, values):
        t = self.torch.tensor(values, dtype=self.torch.float64, device=self.dev)
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t.tolist()]


def measure_gate_batch(env, paramset, B, steps, warmup, check_oracle, cpu_gates_per_thread, no_cpu_baseline, seed=2):
    """K timed steps of EvalBinGate(NAND) over one batch of B gates per GPU for `paramset`; returns (on rank 0) the fields of
    a bench line: value / e2e / roofline / checks / cpu_baseline.  Every rank runs it; only rank 0 gets the dict."""
    import ctypes as C

    import mkfhe_b200 as mk
    from mkfhe_b200 import shard

    torch, dist, dev, rank, world = env.torch, env.dist, env.dev, env.rank, env.world
    method = mk.MKNTRU_LWE if "LWE" in paramset else mk.MKNTRU
    # ---- keys: rank 0 generates on the host (the reference's own keygen when oracle/_ref is present) and loads them; the
    # RESIDENT images are then broadcast device to device into every other rank's context (mkfhe_b200/shard.py) ----
    ref, bits, keys = None, None, None
    if rank == 0:
        P, keys, ct1, b1, ct2, b2, bits, ref = make_workload(paramset, method, B, seed=seed, encrypt=False)
    else:
        P = mk.derive_params(mk.PARAMSETS[paramset], method)
    P.device = env.local
    ctx = mk.BinFHEContext().GenerateFromParams(P)
    ctx.set_stream(env.stream.cuda_stream)
    lwe = ctx.is_lwe
    torch.cuda.synchronize()
    t0 = time.time()
    if rank == 0:
        ctx.LoadKeys(keys["bk"], keys["pkey"], ksk=keys.get("ksk"), kskA=keys.get("kskA"), kskB=keys.get("kskB"),
                     ctnand=keys.get("ctnand"))
    torch.cuda.synchronize()
    t_h2d = time.time() - t0
    env.barrier()
    t0 = time.time()
    key_bytes_total = shard.replicate_key_images(ctx, dev, dist)
    env.barrier()
    t_bcast = time.time() - t0 if world > 1 else 0.0
    t0 = time.time()
    if world > 1:  # the same broadcast once more (idempotent): the steady-state cost, without first-use set-up of the buffers
        shard.replicate_key_images(ctx, dev, dist)
        env.barrier()
    t_bcast2 = time.time() - t0 if world > 1 else 0.0

    # ---- inputs: B distinct gates (rank 0), broadcast; rank r works on them rotated by r ----
    expect, nuniq = None, None
    if rank == 0:
        if ref is not None:
            ct1, b1, ct2, b2, expect, nuniq = distinct_gate_inputs(ctx, P, ref, np.random.default_rng(seed), B)
        pack = [torch.from_numpy(np.ascontiguousarray(v).view(np.int32)).to(dev) for v in (ct1, ct2, b1, b2)]
    else:
        pack = [torch.empty(B * P.k * P.n, dtype=torch.int32, device=dev).view(B, P.k, P.n) for _ in range(2)] + \
               [torch.empty(B, dtype=torch.int32, device=dev) for _ in range(2)]
    if world > 1:
        for t in pack:
            dist.broadcast(t, 0)
        pack = [torch.roll(t, shifts=rank, dims=0).contiguous() for t in pack]
    d_ct1, d_ct2, d_b1, d_b2 = pack
    d_out = torch.empty_like(d_ct1)
    d_bout = torch.empty_like(d_b1)
    h_ct1, h_ct2, h_b1, h_b2 = (t.cpu().pin_memory() for t in pack)
    h_out, h_bout = torch.empty_like(h_ct1).pin_memory(), torch.empty_like(h_b1).pin_memory()
    stream = env.stream

    def step_dev():
        ctx.EvalBinGateDevice(d_ct1, d_ct2, d_out, d_b1 if lwe else None, d_b2 if lwe else None, d_bout if lwe else None)

    hp = lambda t: C.c_void_p(t.data_ptr())

    def step_host():  # the public C-ABI call with HOST buffers: H2D + kernels + D2H, synchronous
        ctx._ck(ctx._L.mkfhe_eval_nand(ctx._h, hp(h_ct1), hp(h_b1) if lwe else None, hp(h_ct2), hp(h_b2) if lwe else None,
                                       hp(h_out), hp(h_bout) if lwe else None, B), "mkfhe_eval_nand")

    for _ in range(warmup):
        step_dev()
    torch.cuda.synchronize()
    launches0 = ctx.launch_count()

    # ---- timed: EXACTLY K steps, device events per step, L2 flushed between steps ----
    sampler = ClockSampler(env.smi_index)
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    rot_ms, ks_ms = [], []
    env.barrier()
    wall0 = time.time()
    for i in range(steps):
        env.flush.fill_(i & 0xff)
        ev[i][0].record(stream)
        step_dev()
        ev[i][1].record(stream)
        a, b = ctx.last_gate_timing()  # waits for the step; kernel-only split for the roofline
        rot_ms.append(a)
        ks_ms.append(b)
    env.barrier()
    wall = time.time() - wall0
    launches = ctx.launch_count() - launches0
    dev_ms = sum(e0.elapsed_time(e1) for e0, e1 in ev)

    # ---- e2e: the same K steps through the host-buffer entry point; with several GPUs every step also gathers the ranks'
    # results on rank 0 (NCCL all_gather of the step's output rows), inside the timed region ----
    step_host()
    gbuf = shard.gather_buffer(d_out, B * world, dist)
    gathered = shard.gather_results(d_out, B * world, dist, out=gbuf)
    env.barrier()
    gather_ms = 0.0
    t0 = time.time()
    for i in range(steps):
        step_host()
        if world > 1:
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g0.record(stream)
            gathered = shard.gather_results(d_out, B * world, dist, out=gbuf)
            g1.record(stream)
            g1.synchronize()
            gather_ms += g0.elapsed_time(g1)
    env.barrier()
    e2e_s = time.time() - t0
    clocks = sampler.stop() if rank == 0 else None
    dev_ms, e2e_ms, rot_sum, ks_sum, gather_ms = env.max_over_ranks([dev_ms, e2e_s * 1e3, sum(rot_ms), sum(ks_ms), gather_ms])

    res = None
    if rank == 0:
        # correctness inside the bench run (outside the timed region)
        out_np = d_out.cpu().numpy().view(np.uint32)
        bo = d_bout.cpu().numpy().view(np.uint32) if lwe else None
        assert (out_np == h_out.numpy().view(np.uint32)).all(), "device-pointer and host-pointer paths disagree"
        assert gathered is None or world == 1 or (gathered[:B].cpu().numpy().view(np.uint32) == out_np).all()
        nand_checked = None
        if ref is not None and expect is not None:  # EVERY output, decrypted on the GPU with the reference's secret key
            ctx.LoadSecretKey(ref.sk())
            got = ctx.Decrypt(out_np, bo).astype(np.int64)
            bad = np.nonzero(got != expect)[0]
            assert len(bad) == 0, f"{len(bad)} of {B} bootstrapped gates do not decrypt to NAND (first: gate {bad[0]})"
            sample = list(range(0, B, max(1, B // 16)))[:16]  # and a sample with the reference's own Decrypt
            assert [ref.decrypt(out_np[i].ravel(), int(bo[i]) if lwe else 0) for i in sample] == [int(expect[i]) for i in sample]
            nand_checked = int(B)
        oracle_checked = check_against_oracle(P, keys, ct1, b1, ct2, b2, out_np, bo, check_oracle)
        wm = work_model(P)
        total_gates = B * world * steps
        value = total_gates / (dev_ms * 1e-3)
        e2e = total_gates / (e2e_ms * 1e-3)
        bfly_peak = ctx.measure_butterfly_peak()
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        rot_s = rot_sum * 1e-3 / steps  # average blind-rotation launch duration (max over ranks)
        mm_ach = wm["modmul"] * B / rot_s  # modular multiplications per second (SURVEY 8d per-gate figure x gates per launch)
        key_ach = wm["key_bytes"] * B / rot_s / 1e9
        traffic = ncu_traffic(paramset, B)
        sm_hz = 1e6 * float((clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz") or 1965.0)
        floor_peak = 148 * 4 * 32 * sm_hz / 8.0  # analytic: IMAD.HI (4) + 2 IMAD (2 + 2) fma-pipe cycles per warp-butterfly
        data = (f"synthetic gate batch: reference keygen + Encrypt (oracle/_ref), {nuniq} distinct gates"
                + (" of refreshed ciphertexts (one NAND level first)" if lwe else " of fresh encryptions")
                if ref is not None else "synthetic (uniform residues)")
        res = {
            "value": value, "ms_per_step": dev_ms / steps, "data": data,
            "config": gates_config(paramset, B, world, nuniq),
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int((2 * B * P.k * P.n + (2 * B if lwe else 0)) * 4),
                    "d2h_bytes_per_step": int((B * P.k * P.n + (B if lwe else 0)) * 4),
                    "gather_s_per_step": gather_ms * 1e-3 / steps,
                    "note": "mkfhe_eval_nand on host buffers per rank; N > 1: + all_gather of the results to rank 0, timed"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "int32", "kernel": "blind_rotate", "achieved": mm_ach / 1e12, "peak": bfly_peak / 1e12,
                         "unit": "Tmodmul/s", "frac": mm_ach / bfly_peak, "traffic": traffic,
                         "frac_of_analytic_floor": mm_ach / floor_peak, "analytic_floor_peak": floor_peak / 1e12,
                         "model": "achieved = SURVEY 8d modular multiplications per gate (butterflies + pointwise MACs) x gates "
                                  "per launch / launch time; peak = register-resident lazy Shoup butterflies/s measured on this "
                                  "GPU in this run (mkfhe_measure_butterfly_peak: nothing but butterflies, no memory traffic); "
                                  "analytic floor = 8 fma-pipe cycles per warp-butterfly at the sampled SM clock",
                         "modmul_per_gate": wm["modmul"], "kernel_ms": rot_s * 1e3, "share_of_step": rot_sum / dev_ms},
            "roofline_hbm": {"bound": "hbm", "kernel": "blind_rotate key stream", "achieved": key_ach, "peak": hbm_peak,
                             "unit": "GB/s", "frac": key_ach / hbm_peak, "traffic": traffic,
                             "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback 6650",
                             "bytes_per_gate": wm["key_bytes"], "note": "algorithmic key bytes if every gate streamed its own copy"},
            "kernels_ms": {"blind_rotate": rot_sum / steps, "key_switch": ks_sum / steps},
            "clocks": clocks, "wall_s_timed": wall, "key_bytes_replicated": key_bytes_total,
            "key_replication": {"nccl_init_s": env.nccl_init_s, "h2d_and_relayout_s": t_h2d, "bcast_s": t_bcast,
                                "bcast_repeat_s": t_bcast2, "bytes": key_bytes_total, "what": "rank 0: pageable H2D + Montgomery/register-order/u16 re-layout; "
                                "then the resident images broadcast device to device (NCCL)"},
            "key_broadcast_s": t_bcast, "gather_s": gather_ms * 1e-3 / steps,
            "nand_checked": nand_checked, "oracle_checked": oracle_checked,
        }
        if world == 1 and not no_cpu_baseline:
            cores = os.cpu_count() or 1
            gates = cores * cpu_gates_per_thread  # bounded sample: ~10-20 s of host work
            # with synthetic keys (the party sweep: the reference's keygen takes minutes at k >= 8) the CPU leg is the C port
            # of the path on the same synthetic keys' shapes
            gps, kind, secs = cpu_gates_per_s(paramset, cores, gates, ref, force_port=ref is None and bool(os.environ.get("MKFHE_BENCH_SYNTHETIC_KEYS")))
            res["cpu_baseline"] = {"value": gps, "unit": UNIT, "cores": cores, "kind": kind,
                                   "sample": f"{gates} EvalBinGate(NAND) on {cores} threads in {secs:.1f}s, same paramset"}
            if ref is not None:  # the reference example's own measurement: one gate, one thread (boolean-mkntru.cpp:34-38)
                ms = sorted(ref.time_gate(5))
                res["cpu_baseline"]["single_thread_ms"] = ms[len(ms) // 2]
                res["cpu_baseline"]["single_thread_sample"] = "median of 5 single EvalBinGate(NAND) calls, steady_clock, 1 thread"
    if rank == 0:
        # BASELINE.json configs[0] / the reference examples' own measurement: ONE gate per call (the drop-in accumulator's case).
        # Device-resident kernel time and the host-buffer C-ABI call, median of 5 after 2 warm-up calls; small batches run on a
        # thread-block cluster of `sms_per_gate` SMs (mkfhe_b200/csrc/cluster_kernels.cuh)
        one = [t[:1] for t in (d_ct1, d_ct2, d_out, d_b1, d_b2, d_bout)]
        lat, lat_host = [], []
        for _ in range(7):
            ctx.EvalBinGateDevice(one[0], one[1], one[2], one[3] if lwe else None, one[4] if lwe else None, one[5] if lwe else None)
            a, b = ctx.last_gate_timing()
            lat.append(a + b)
        sms_per_gate = ctx.last_gate_sms()
        for _ in range(7):
            t0 = time.perf_counter()
            ctx._ck(ctx._L.mkfhe_eval_nand(ctx._h, hp(h_ct1), hp(h_b1) if lwe else None, hp(h_ct2), hp(h_b2) if lwe else None,
                                           hp(h_out), hp(h_bout) if lwe else None, 1), "mkfhe_eval_nand")
            lat_host.append((time.perf_counter() - t0) * 1e3)
        res["single_gate"] = {"ms": sorted(lat[2:])[2], "e2e_ms": sorted(lat_host[2:])[2], "sms_per_gate": sms_per_gate,
                              "what": "one EvalBinGate(NAND) per call: kernel time (CUDA events) and the synchronous host-buffer call"}
    ctx.close()
    del ctx
    torch.cuda.empty_cache()
    return res


def run_gpu_arm(args):
    env = GpuEnv()
    main = measure_gate_batch(env, args.paramset, args.batch, args.steps, args.warmup, args.check_oracle,
                              args.cpu_baseline_gates_per_thread, args.no_cpu_baseline)
    also = None
    if args.paramset == "STD100_MKNTRU_LWE" and not args.no_also:
        # north_star names mk-acc-xzw.cpp: the boolean-mkntru (XZW, ternary) shape at the same k and batch, as its own block
        also = measure_gate_batch(env, "STD100_MKNTRU", args.batch, args.also_steps, min(args.warmup, 3), args.check_oracle,
                                  max(1, args.cpu_baseline_gates_per_thread // 2), args.no_cpu_baseline, seed=3)
    if env.rank == 0:
        line = {"metric": METRIC, "value": main["value"], "unit": UNIT, "n_gpus": env.world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u32"}
        line.update({k: v for k, v in main.items() if k not in ("value", "ms_per_step")})
        if also is not None:
            also.update({"metric": METRIC, "unit": UNIT, "steps": args.also_steps, "n_gpus": env.world})
            line["also"] = {"boolean_mkntru": also}
        emit(line)
    if env.dist is not None:
        env.dist.barrier()
        env.dist.destroy_process_group()


# ---------------------------------------------------------------- configs[3]: a fixed job sharded over the GPUs
def run_sharded_arm(args):
    """--workload sharded: --total-gates NAND gates (default paramset STD100_MKNTRU_2, k = 4) split into contiguous shards,
    one per GPU; keys loaded on rank 0 and replicated as resident images (NCCL broadcast, timed); inputs counter-based
    (mkfhe_b200/shard.py) and generated on the device by the rank that owns them; one step = every rank bootstraps its whole
    shard in launches of --chunk gates AND the results are gathered on rank 0 (timed, inside the wall clock: `gather_s`).
    Strong scaling: the job is fixed.  The XOR digest of all outputs does not depend on the number of GPUs; --check-oracle
    gates spread over the whole job are regenerated on the CPU from their counters and re-computed by the oracle."""
    import mkfhe_b200 as mk
    from mkfhe_b200 import shard

    env = GpuEnv()
    torch, dist, dev, rank, world, stream = env.torch, env.dist, env.dev, env.rank, env.world, env.stream
    method = mk.MKNTRU_LWE if "LWE" in args.paramset else mk.MKNTRU
    P = mk.derive_params(mk.PARAMSETS[args.paramset], method)
    keys = None
    if rank == 0:
        os.environ["MKFHE_BENCH_SYNTHETIC_KEYS"] = "1"  # uniform key residues: the job is about sharding and throughput
        _, keys, *_ = make_workload(args.paramset, method, 1, seed=4)
    P.device = env.local
    ctx = mk.BinFHEContext().GenerateFromParams(P)
    ctx.set_stream(stream.cuda_stream)
    lwe = ctx.is_lwe
    torch.cuda.synchronize()
    t0 = time.time()
    if rank == 0:
        ctx.LoadKeys(keys["bk"], keys["pkey"], ksk=keys.get("ksk"), kskA=keys.get("kskA"), kskB=keys.get("kskB"), ctnand=keys.get("ctnand"))
    torch.cuda.synchronize()
    t_h2d = time.time() - t0
    env.barrier()
    t0 = time.time()
    key_bytes_total = shard.replicate_key_images(ctx, dev, dist)
    env.barrier()
    t_bcast = time.time() - t0 if world > 1 else 0.0
    t0 = time.time()
    if world > 1:
        shard.replicate_key_images(ctx, dev, dist)
        env.barrier()
    t_bcast2 = time.time() - t0 if world > 1 else 0.0
    T, seed = args.total_gates, 4
    lo, hi = shard.shard_range(T, world, rank)
    chunks = [(a, min(hi, a + args.chunk)) for a in range(lo, hi, args.chunk)]
    mine = torch.empty((hi - lo, P.k, P.n), dtype=torch.int32, device=dev)  # this rank's results, resident until the gather
    mine_b = torch.empty(hi - lo, dtype=torch.int32, device=dev) if lwe else None

    def run_shard():
        ms = 0.0
        for a, b in chunks:
            inp = shard.gate_inputs_torch(P, seed, a, b, dev)  # not timed by the events: input synthesis is not part of the path
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            ctx.EvalBinGateDevice(inp["ct1"], inp["ct2"], mine[a - lo:b - lo], inp.get("b1"), inp.get("b2"),
                                  mine_b[a - lo:b - lo] if lwe else None)
            e1.record(stream)
            e1.synchronize()
            ms += e0.elapsed_time(e1)
        return ms

    for _ in range(args.warmup):  # warm-up on the first chunk only (the job itself is minutes long)
        a, b = chunks[0]
        inp = shard.gate_inputs_torch(P, seed, a, min(b, a + 592), dev)
        ctx.EvalBinGateDevice(inp["ct1"], inp["ct2"], torch.empty_like(inp["ct1"]), inp.get("b1"), inp.get("b2"),
                              torch.empty(inp["ct1"].shape[0], dtype=torch.int32, device=dev) if lwe else None)
    torch.cuda.synchronize()
    sampler = ClockSampler(env.smi_index)
    if rank == 0:
        sampler.start()
    launches0 = ctx.launch_count()
    total_ms, gather_ms, gathered = 0.0, 0.0, None
    gbuf, gbuf_b = shard.gather_buffer(mine, T, dist), (shard.gather_buffer(mine_b, T, dist) if lwe else None)
    env.barrier()
    wall0 = time.time()
    for step in range(args.steps):
        total_ms += run_shard()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record(stream)
        gathered = shard.gather_results(mine, T, dist, out=gbuf)  # rank 0 ends up with all T output rows, in gate order
        gathered_b = shard.gather_results(mine_b, T, dist, out=gbuf_b) if lwe else None
        g1.record(stream)
        g1.synchronize()
        gather_ms += g0.elapsed_time(g1)
    env.barrier()
    wall = time.time() - wall0
    launches = ctx.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    total_ms, gather_ms = env.max_over_ranks([total_ms, gather_ms])
    if rank == 0:
        h = shard.xor_hash_torch(gathered, 0)                   # the digest is taken over the GATHERED table, on the device
        checked = 0
        if args.check_oracle:  # gates spread over the whole job (every shard), regenerated on the CPU from their counters
            idx = sorted(set(int(x) for x in np.linspace(0, T - 1, args.check_oracle)))
            g = shard.gate_inputs_np(P, seed, idx)
            ti = torch.tensor(idx, dtype=torch.int64, device=dev)
            sel = gathered[ti].cpu().numpy().view(np.uint32)
            selb = gathered_b[ti].cpu().numpy().view(np.uint32) if lwe else None
            checked = check_against_oracle(P, keys, g["ct1"], g.get("b1"), g["ct2"], g.get("b2"), sel, selb, len(idx))
        wm = work_model(P)
        step_s = (total_ms + gather_ms) * 1e-3 / args.steps
        value = T / step_s
        bfly_peak = ctx.measure_butterfly_peak()
        mm = wm["modmul"] * T * args.steps / (total_ms * 1e-3) / world  # per GPU, kernels only
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "u32", "data": "synthetic: uniform key residues, counter-based uniform ciphertexts (seed 4 ^ gate index)",
                "config": {"workload": f"{args.paramset} k={P.k}, {T} independent NAND bootstraps sharded over {world} GPU(s)",
                           "paramset": args.paramset, "total_gates": T, "chunk": args.chunk,
                           "l2": "inputs of one step (>= 147 MB per launch) exceed L2; keys are re-streamed per launch",
                           "parallelism": f"gates sharded x{world}, keys replicated"},
                "gpu_launches": int(launches), "output_xor_digest": f"{h:016x}", "oracle_checked": checked,
                "compute_s": total_ms * 1e-3 / args.steps, "gather_s": gather_ms * 1e-3 / args.steps,
                "value_compute_only": T * args.steps / (total_ms * 1e-3),
                "roofline": {"bound": "int32", "kernel": "blind_rotate", "achieved": mm / 1e12, "peak": bfly_peak / 1e12,
                             "unit": "Tmodmul/s per GPU", "frac": mm / bfly_peak, "traffic": None, "modmul_per_gate": wm["modmul"]},
                "clocks": clocks, "wall_s_timed": wall, "key_bytes_replicated": key_bytes_total,
                "key_replication": {"nccl_init_s": env.nccl_init_s, "h2d_and_relayout_s": t_h2d, "bcast_s": t_bcast, "bcast_repeat_s": t_bcast2,
                                    "bytes": key_bytes_total},
                "key_broadcast_s": t_bcast}
        emit(line)
    ctx.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


# ---------------------------------------------------------------- configs[4]: batched transform micro-benchmark
def run_ntt_arm(args):
    """--workload ntt: NativePoly::SetFormat on --ntt-batch polynomials of --ntt-n coefficients (forward, then inverse), in
    place in HBM.  Both loaders are measured in the same run: the bulk-copy (TMA) ring that is the default and the
    register-prefetch kernel it replaced (MKFHE_NTT_VARIANT=ldg).  Every variant: K timed forward + inverse pairs, K raised
    until the timed region lasts >= --min-seconds so that the clock sampler sees the GPU under load; parity of 1024 sampled
    polynomials against the CPU oracle in both directions and the round-trip identity on all, outside the timed region."""
    import torch

    import mkfhe_b200 as mk
    from mkfhe_b200 import context as mctx

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the GPU arm has no CPU fallback")
    dev = torch.device("cuda", 0)
    N, B = args.ntt_n, args.ntt_batch
    Q = mctx.previous_prime(mctx.first_prime(27, 2 * N), 2 * N)
    psi = mctx.root_of_unity(2 * N, Q)
    P = mk.derive_params(mk.PARAMSETS["STD100_MKNTRU"], mk.MKNTRU)
    P.N, P.Q, P.psi, P.device = N, Q, psi, 0
    ctx = mk.BinFHEContext().GenerateFromParams(P)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    gen = torch.Generator(device=dev)
    gen.manual_seed(5)
    x = torch.randint(0, Q, (B, N), dtype=torch.int32, device=dev, generator=gen)
    x0 = x.clone()
    o = make_oracle(P)
    samp = x0[:: max(1, B // 1024)][:1024].contiguous()
    samp_np = samp.cpu().numpy().view(np.uint32)
    want_f, want_i = o.ntt(samp_np, inverse=False), o.ntt(samp_np, inverse=True)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    bfly_peak = ctx.measure_butterfly_peak()
    bytes_per = 2 * N * 4
    bfly = (N // 2) * (N.bit_length() - 1)

    def measure(variant):
        if variant == "tma":
            os.environ.pop("MKFHE_NTT_VARIANT", None)
        else:
            os.environ["MKFHE_NTT_VARIANT"] = variant
        for _ in range(args.warmup):
            ctx.NTTDevice(x, inverse=False)
            ctx.NTTDevice(x, inverse=True)
        torch.cuda.synchronize()
        assert torch.equal(x, x0), f"{variant}: round trip is not the identity"
        for inv, want in ((False, want_f), (True, want_i)):
            y = samp.clone()
            ctx.NTTDevice(y, inverse=inv)
            torch.cuda.synchronize()
            assert (y.cpu().numpy().view(np.uint32) == want).all(), f"{variant}: transform differs from the oracle (inverse={inv})"
        sampler = ClockSampler(0)
        sampler.start()
        launches0 = ctx.launch_count()
        fwd_ms, inv_ms = [], []
        t0 = time.time()
        while len(fwd_ms) < args.steps or time.time() - t0 < args.min_seconds:
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            e[0].record(stream)
            ctx.NTTDevice(x, inverse=False)
            e[1].record(stream)
            ctx.NTTDevice(x, inverse=True)
            e[2].record(stream)
            e[2].synchronize()
            fwd_ms.append(e[0].elapsed_time(e[1]))
            inv_ms.append(e[1].elapsed_time(e[2]))
        launches = ctx.launch_count() - launches0
        clocks = sampler.stop()
        assert torch.equal(x, x0), f"{variant}: round trip is not the identity"
        tf, ti = sum(fwd_ms) / len(fwd_ms) * 1e-3, sum(inv_ms) / len(inv_ms) * 1e-3
        return {"value": 2 * B / (tf + ti), "steps": len(fwd_ms), "ms_per_step": (tf + ti) * 1e3, "forward_ms": tf * 1e3,
                "inverse_ms": ti * 1e3, "best_forward_ms": min(fwd_ms), "best_inverse_ms": min(inv_ms), "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "kernel": f"ntt_batch ({variant} loader, forward)", "achieved": B * bytes_per / tf / 1e9,
                             "peak": hbm_peak, "unit": "GB/s", "frac": B * bytes_per / tf / 1e9 / hbm_peak, "traffic": None,
                             "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback 6650"},
                "roofline_inverse": {"bound": "hbm", "achieved": B * bytes_per / ti / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                     "frac": B * bytes_per / ti / 1e9 / hbm_peak},
                "roofline_int32": {"achieved": B * bfly / tf / 1e12, "peak": bfly_peak / 1e12, "unit": "Tbutterfly/s",
                                   "frac": B * bfly / tf / bfly_peak, "frac_inverse": B * bfly / ti / bfly_peak,
                                   "model": "(N/2) log2 N butterflies per transform against the register-resident butterfly rate "
                                            "measured in this run; at N = 2048 this ceiling (peak / 11264 transforms/s) and the HBM "
                                            "ceiling (16 KiB per transform) are within 10% of each other"},
                "clocks": clocks}

    res = {v: measure(v) for v in args.ntt_variants.split(",")}
    os.environ.pop("MKFHE_NTT_VARIANT", None)
    main = res.get("tma") or next(iter(res.values()))
    line = {"metric": "negacyclic NTT transforms/sec", "value": main["value"], "unit": "transforms/s", "n_gpus": 1,
            "steps": main["steps"], "warmup": args.warmup, "ms_per_step": main["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic: uniform residues (torch philox, seed 5)",
            "config": {"workload": f"NTT + INTT micro-benchmark, N={N}, Q={Q}, batch {B} polynomials in HBM",
                       "l2": f"working set {B * N * 4 >> 20} MiB exceeds L2"},
            "oracle_checked": int(len(samp)) * 2}
    line.update({k: v for k, v in main.items() if k not in ("value", "steps", "ms_per_step")})
    line["variants"] = res
    emit(line)
    ctx.close()


def run_keygen_arm(args):
    """--workload keygen: bootstrapping-key generation for --paramset, randomness supplied (synthetic small
    residues).  value = evks/s with inputs and output resident in HBM; e2e = the same through host buffers (H2D of the
    randomness, D2H of the key).  cpu_baseline = the oracle's restatement of KeyGenXZW (one thread, bounded sample); the
    reference's own KeyGenAcc for k = 2 takes 2.46 s on one core."""
    import ctypes as C

    import torch

    import mkfhe_b200 as mk

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the GPU arm has no CPU fallback")
    dev = torch.device("cuda", 0)
    method = mk.MKNTRU_LWE if "LWE" in args.paramset else mk.MKNTRU
    P = mk.derive_params(mk.PARAMSETS[args.paramset], method)
    P.device = 0
    ctx = mk.BinFHEContext().GenerateFromParams(P)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    ell, N, Q = P.digitsG - 1, P.N, P.Q
    S = 2 if P.variant == 0 else 1
    E = P.k * S * P.n + 1
    rng = np.random.default_rng(8)
    small = lambda shape: (rng.integers(-3, 4, size=shape) % Q).astype(np.uint32)
    sinv = rng.integers(1, Q, size=(P.k, N), dtype=np.uint32)
    crs = rng.integers(0, Q, size=(ell, N), dtype=np.uint32)
    r, e = small((E, N)), small((E, ell, 2, N))
    m = rng.integers(0, 2, size=E, dtype=np.uint8)
    out = np.empty((E, ell, 2, N), dtype=np.uint32)
    t = lambda a: torch.from_numpy(a.view(np.int32) if a.dtype == np.uint32 else a).to(dev)
    d_sinv, d_crs, d_r, d_e, d_m = t(sinv), t(crs), t(r), t(e), t(m)
    d_out = torch.empty((E, ell, 2, N), dtype=torch.int32, device=dev)
    L, h = ctx._L, ctx._h
    vp = lambda x: C.c_void_p(x.data_ptr() if hasattr(x, "data_ptr") else x.ctypes.data)

    def dev_call():
        ctx._ck(L.mkfhe_keygen_bk(h, vp(d_sinv), vp(d_crs), vp(d_r), vp(d_e), vp(d_m), vp(d_out), 0), "mkfhe_keygen_bk")

    def host_call():
        ctx._ck(L.mkfhe_keygen_bk(h, vp(sinv), vp(crs), vp(r), vp(e), vp(m), vp(out), 0), "mkfhe_keygen_bk")

    for _ in range(args.warmup):
        dev_call()
    sampler = ClockSampler(0)
    sampler.start()
    launches0 = ctx.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    torch.cuda.synchronize()
    ev[0].record(stream)
    for _ in range(args.steps):
        dev_call()
    ev[1].record(stream)
    ev[1].synchronize()
    dev_ms = ev[0].elapsed_time(ev[1])
    launches = ctx.launch_count() - launches0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        host_call()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop()
    assert (d_out.cpu().numpy().view(np.uint32) == out).all()
    # parity sample + CPU baseline: the oracle's restatement on one thread
    o = make_oracle(P)
    idx = [0, 1, E // 2, E - 1]
    for i in idx:
        u = 0 if i == E - 1 else i // (S * P.n)
        kdm = i == E - 1 or (u == 0 and (i % P.n) == 0)
        assert (out[i] == o.keygen_evk(sinv[u], crs, r[i], e[i], int(m[i]), kdm)).all(), "GPU evk differs from the oracle"
    nb = min(E, 256)
    t0 = time.perf_counter()
    for i in range(nb):
        o.keygen_evk(sinv[0], crs, r[i], e[i], int(m[i]), False)
    cpu_s = time.perf_counter() - t0
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    step_s = dev_ms * 1e-3 / args.steps
    bytes_step = E * (N + 2 * ell * N + 2 * ell * N) * 4  # r + e read, evk written (device-resident call also copies e, r once)
    line = {"metric": "uni-encryption evaluation keys generated/sec", "value": E / step_s, "unit": "evks/s", "n_gpus": 1,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic randomness (small residues), uniform secrets / CRS",
            "config": {"workload": f"{args.paramset} bootstrapping-key generation, {E} evks of {2 * ell * N * 4 >> 10} KiB",
                       "paramset": args.paramset, "l2": f"key of {E * 2 * ell * N * 4 >> 20} MiB written per step"},
            "e2e": {"value": E * args.steps / e2e_s, "unit": "evks/s", "h2d_bytes_per_step": int(E * (N + 2 * ell * N) * 4),
                    "d2h_bytes_per_step": int(E * 2 * ell * N * 4)},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "keygen_bk_kernel (+ staging copies)", "achieved": bytes_step / step_s / 1e9,
                         "peak": hbm_peak, "unit": "GB/s", "frac": bytes_step / step_s / 1e9 / hbm_peak, "traffic": None,
                         "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback 6650"},
            "cpu_baseline": {"value": nb / cpu_s, "unit": "evks/s", "cores": 1, "kind": "port",
                             "sample": f"{nb} evks through oracle/mkfhe_oracle.c orc_keygen_evk in {cpu_s:.2f}s"},
            "oracle_checked": len(idx), "clocks": clocks}
    emit(line)
    ctx.close()


def run_circuit_arm(args):
    """--workload circuit: --adders independent ripple-carry adders of --adder-bits bits as ONE NAND netlist through
    mkfhe_eval_circuit: the library levelises it and runs one gate batch per level.  value = gates/s with
    the wire table resident in HBM (mkfhe_eval_circuit_dev), e2e = the same through host buffers (whole wire table H2D and
    D2H).  With oracle/_ref present the inputs are the reference's encryptions, a sample of sums is decrypted by the
    reference's Decrypt and the first XOR is re-computed by the reference's EvalBinGate."""
    import torch

    import mkfhe_b200 as mk
    from mkfhe_b200.netlist import ripple_carry_adders

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the GPU arm has no CPU fallback")
    dev = torch.device("cuda", 0)
    method = mk.MKNTRU_LWE if "LWE" in args.paramset else mk.MKNTRU
    bits, inst = args.adder_bits, args.adders
    P, keys, ct1, b1, ct2, b2, mbits, ref = make_workload(args.paramset, method, 64, seed=6)
    P.device = 0
    ctx = mk.BinFHEContext().GenerateFromParams(P)
    lwe = ctx.is_lwe
    ctx.LoadKeys(keys["bk"], keys["pkey"], ksk=keys.get("ksk"), kskA=keys.get("kskA"), kskB=keys.get("kskB"), ctnand=keys.get("ctnand"))
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    nl, sums = ripple_carry_adders(bits, inst)
    gates, W, per = nl.array(), nl.n_wires, 2 * bits + 1
    G = len(gates)
    wires = np.zeros((W, P.k, P.n), dtype=np.uint32)
    wb = np.zeros(W, dtype=np.uint32)
    nu = len(ct1)
    zero = ref.encrypt(0) if ref is not None else (ct1[0].ravel(), int(b1[0]))
    avals, bvals = np.zeros(inst, dtype=np.int64), np.zeros(inst, dtype=np.int64)
    for t in range(inst):
        for i in range(bits):
            j = (t * bits + i) % nu
            wires[per * t + i], wb[per * t + i] = ct1[j], b1[j]
            wires[per * t + bits + i], wb[per * t + bits + i] = ct2[j], b2[j]
            if mbits is not None:
                avals[t] |= int(mbits[0][j]) << i
                bvals[t] |= int(mbits[1][j]) << i
        wires[per * t + 2 * bits] = np.asarray(zero[0], dtype=np.uint32).reshape(P.k, P.n)
        wb[per * t + 2 * bits] = zero[1]
    d_w = torch.from_numpy(wires.view(np.int32)).to(dev)
    d_b = torch.from_numpy(wb.view(np.int32)).to(dev) if lwe else None
    levels = 0

    def step_dev():
        nonlocal levels
        levels = ctx.EvalCircuitDevice(d_w, gates, d_b)

    for _ in range(args.warmup):
        step_dev()
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    sampler.start()
    launches0 = ctx.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record(stream)
    for _ in range(args.steps):
        step_dev()
    ev[1].record(stream)
    ev[1].synchronize()
    dev_ms = ev[0].elapsed_time(ev[1])
    launches = ctx.launch_count() - launches0
    out, ob, _ = ctx.EvalCircuit(wires, gates, wb if lwe else None)  # warm the host path
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out, ob, _ = ctx.EvalCircuit(wires, gates, wb if lwe else None)
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop()
    assert (d_w.cpu().numpy().view(np.uint32).reshape(out.shape) == out).all(), "device-pointer and host-pointer paths disagree"
    if ob is None:
        ob = np.zeros(W, dtype=np.uint32)
    checked = ref_gates = None
    if ref is not None:
        if lwe:  # the reference's MKNTRU gate outputs do not compose, so only MK-LWE sums are meaningful
            idx = list(range(0, inst, max(1, inst // 8)))[:8]
            for t in idx:
                got = sum(ref.decrypt(out[w].ravel(), int(ob[w])) << i for i, w in enumerate(sums[t]))
                assert got == int(avals[t] + bvals[t]), "an adder does not decrypt to a + b"
            checked = len(idx)
        cur = {int(g[j]): (wires[g[j]].ravel(), int(wb[g[j]])) for g in gates[:2] for j in (0, 1) if g[j] < per * inst}
        for i1, i2, o in gates[:4]:
            cur[int(o)] = ref.eval_gate(cur[int(i1)][0], cur[int(i1)][1], cur[int(i2)][0], cur[int(i2)][1])
            assert (out[o].ravel() == cur[int(o)][0]).all() and int(ob[o]) == int(cur[int(o)][1]), "circuit differs from the reference chain"
        ref_gates = 4
    wm = work_model(P)
    step_s = dev_ms * 1e-3 / args.steps
    bfly_peak = ctx.measure_butterfly_peak()
    mm_ach = wm["modmul"] * G / step_s
    line = {"metric": METRIC, "value": G / step_s, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32",
            "data": "synthetic circuit inputs: reference keygen + Encrypt (oracle/_ref)" if ref is not None else "synthetic (uniform residues)",
            "config": {"workload": f"{args.paramset}: {inst} ripple-carry adders of {bits} bits as one NAND netlist ({G} gates, {levels} levels, "
                                   f"{W} wires) through mkfhe_eval_circuit", "paramset": args.paramset, "gates": G, "levels": int(levels),
                       "l2": f"wire table of {W * P.k * P.n * 4 >> 20} MiB, keys of {sum(np.asarray(v).size for v in keys.values()) * 4 >> 20} MiB per level"},
            "e2e": {"value": G * args.steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(W * (P.k * P.n + (1 if lwe else 0)) * 4),
                    "d2h_bytes_per_step": int(W * (P.k * P.n + (1 if lwe else 0)) * 4)},
            "gpu_launches": int(launches),
            "roofline": {"bound": "int32", "kernel": "blind_rotate (one launch per level)", "achieved": mm_ach / 1e12,
                         "peak": bfly_peak / 1e12, "unit": "Tmodmul/s", "frac": mm_ach / bfly_peak, "traffic": None,
                         "model": "as --workload gates, over the whole step (all levels, gather/scatter included)",
                         "modmul_per_gate": wm["modmul"]},
            "clocks": clocks, "adders_checked": checked, "reference_gates_checked": ref_gates}
    if not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        ng = cores * args.cpu_baseline_gates_per_thread
        gps, kind, secs = cpu_gates_per_s(args.paramset, cores, ng, ref)
        line["cpu_baseline"] = {"value": gps, "unit": UNIT, "cores": cores, "kind": kind,
                                "sample": f"{ng} EvalBinGate(NAND) on {cores} threads in {secs:.1f}s, same paramset (the reference "
                                          "runs a circuit as a loop of such calls)"}
    emit(line)
    ctx.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--paramset", default="STD100_MKNTRU_LWE")
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--ref-gates-per-thread", type=int, default=8, help="--impl reference: gates per host thread per step")
    ap.add_argument("--cpu-baseline-gates-per-thread", type=int, default=32)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--synthetic-keys", action="store_true", help="uniform random key residues instead of the reference keygen "
                    "(throughput does not depend on key values; the reference's keygen takes minutes at k >= 8)")
    ap.add_argument("--check-oracle", type=int, default=64, help="gates to re-compute on the CPU oracle after the run (all host threads)")
    ap.add_argument("--no-also", action="store_true", help="default workload only: skip the boolean-mkntru (XZW) block")
    ap.add_argument("--also-steps", type=int, default=5)
    ap.add_argument("--workload", default="gates", choices=["gates", "sharded", "ntt", "keygen", "circuit"],
                    help="gates: configs[1]/[2] (batch per GPU); sharded: configs[3] (--total-gates split over the GPUs, "
                         "counter-based inputs); ntt: configs[4] (batched transform micro-benchmark)")
    ap.add_argument("--total-gates", type=int, default=1 << 17)
    ap.add_argument("--chunk", type=int, default=16384, help="sharded: gates per launch")
    ap.add_argument("--adders", type=int, default=512, help="circuit: independent adders in the netlist")
    ap.add_argument("--adder-bits", type=int, default=8)
    ap.add_argument("--ntt-n", type=int, default=2048)
    ap.add_argument("--ntt-batch", type=int, default=1 << 18)
    ap.add_argument("--ntt-variants", default="tma,1,ldg", help="loaders to measure: tma (the default kernel: bulk-copy ring, two "
                    "polynomials per slot), 1 (one polynomial per slot), b (1 with the Barrett-style final fold), ldg (register prefetch)")
    ap.add_argument("--min-seconds", type=float, default=2.5, help="ntt: minimum length of a timed region (clock sampling)")
    args = ap.parse_args()
    if args.synthetic_keys:
        os.environ["MKFHE_BENCH_SYNTHETIC_KEYS"] = "1"
    capture_stdout()
    if args.impl == "reference":
        run_reference_arm(args)
    elif args.workload == "sharded":
        run_sharded_arm(args)
    elif args.workload == "ntt":
        run_ntt_arm(args)
    elif args.workload == "keygen":
        run_keygen_arm(args)
    elif args.workload == "circuit":
        run_circuit_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
