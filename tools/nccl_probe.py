#!/usr/bin/env python
"""What the collectives of this box cost: broadcast / gather / all_reduce of 64 MiB .. 1 GiB between the GPUs of one node, timed
with CUDA events after a warm-up, plus the transports NCCL chose (NCCL_DEBUG=INFO lines).  Run under torchrun:
    NCCL_DEBUG=INFO python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/nccl_probe.py"""
import json
import os
import time

import torch
import torch.distributed as dist

rank, world, local = (int(os.environ[v]) for v in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
res = {"world": world, "p2p": [[torch.cuda.can_device_access_peer(i, j) if i != j else None for j in range(world)] for i in range(world)] if rank == 0 else None}


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    dist.barrier()
    out = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        out.append(a.elapsed_time(b))
    t = torch.tensor([min(out)], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


for mb in (64, 256, 1024):
    x = torch.empty(mb << 18, dtype=torch.int32, device=dev)
    ms = timed(lambda: dist.broadcast(x, 0))
    res[f"broadcast_{mb}MiB_ms"] = ms
    res[f"broadcast_{mb}MiB_GBs"] = (mb / 1024 * 1.073741824) / (ms * 1e-3)
    bufs = [torch.empty_like(x) for _ in range(world)] if rank == 0 else None
    ms = timed(lambda: dist.gather(x, bufs, dst=0))
    res[f"gather_{mb}MiB_per_rank_ms"] = ms
    res[f"gather_{mb}MiB_into_rank0_GBs"] = ((world - 1) * mb / 1024 * 1.073741824) / (ms * 1e-3)
    del bufs
    ms = timed(lambda: dist.all_reduce(x))
    res[f"all_reduce_{mb}MiB_ms"] = ms
    del x
if rank == 0:
    print(json.dumps(res))
dist.barrier()
dist.destroy_process_group()
