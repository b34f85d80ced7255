"""Multi-GPU host logic on CPU: world size 2 over gloo (BASELINE.json configs[3]).

No ciphertext arithmetic happens here (there is no CPU fallback to call): the "gate" of these tests is a stand-in that only
depends on the gate's own inputs and the replicated keys, which is all the sharding logic may assume."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mkfhe_b200 as mk
from mkfhe_b200 import shard


def test_shard_ranges_cover_and_are_contiguous():
    for total in (0, 1, 7, 4096, 1 << 20):
        for world in (1, 2, 3, 4, 8):
            r = [shard.shard_range(total, world, k) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            sizes = [hi - lo for lo, hi in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.shard_range(4, 2, 2)


@pytest.mark.parametrize("paramset,method", [("STD100_MKNTRU_2", mk.MKNTRU), ("STD100_MKNTRU_LWE", mk.MKNTRU_LWE)])
def test_counter_inputs_host_and_device_generators_agree(paramset, method):
    P = mk.derive_params(mk.PARAMSETS[paramset], method)
    a = shard.gate_inputs_np(P, 4, np.arange(1000, 1037))
    b = shard.gate_inputs_torch(P, 4, 1000, 1037, "cpu")
    assert set(a) == set(b)
    for nme in a:
        assert (a[nme] == b[nme].numpy().view(np.uint32)).all()
        assert a[nme].max() < P.q
    # a single gate regenerated on its own is the same gate
    one = shard.gate_inputs_np(P, 4, [1020])
    assert (one["ct1"][0] == a["ct1"][20]).all() and (one["ct2"][0] == a["ct2"][20]).all()
    # streams, seeds and gates are independent
    assert (a["ct1"] != a["ct2"]).mean() > 0.99
    assert (shard.gate_inputs_np(P, 5, [1020])["ct1"] != one["ct1"]).mean() > 0.99
    # residues are spread over [0, q)
    assert abs(a["ct1"].mean() / P.q - 0.5) < 0.01


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_gate(P, keys, inp):
    """stand-in for the GPU gate: depends on the gate's inputs and on the replicated keys only"""
    mixk = int(keys["bk"].to(torch.int64).sum().item() + keys["ctnand"].to(torch.int64).sum().item()) % P.q
    return ((inp["ct1"].to(torch.int64) * 3 + inp["ct2"].to(torch.int64) + mixk) % P.q).to(torch.int32)


def _worker(rank, world, port, total, out_path):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    P = mk.derive_params(mk.PARAMSETS["STD100_MKNTRU_2"], mk.MKNTRU)
    keys = None
    if rank == 0:
        rng = np.random.default_rng(11)
        keys = {"bk": rng.integers(0, P.Q, 5000, dtype=np.uint32), "pkey": rng.integers(0, P.Q, 300, dtype=np.uint32),
                "ksk": rng.integers(0, P.qKS, 777, dtype=np.uint32), "ctnand": rng.integers(0, P.q, P.k * P.n, dtype=np.uint32)}
    dkeys = shard.broadcast_keys(keys, "cpu", dist)
    lo, hi = shard.shard_range(total, world, rank)
    inp = shard.gate_inputs_torch(P, 4, lo, hi, "cpu")
    out = _fake_gate(P, dkeys, inp)
    part = shard.xor_hash_np(out.numpy().view(np.uint32), lo)
    parts = [None] * world
    dist.all_gather_object(parts, part)
    allout = shard.gather_results(out, total, dist)
    if rank == 0:
        h = 0
        for p in parts:
            h ^= p
        assert shard.xor_hash_torch(allout, 0, block=16) == h  # the device-side digest of the gathered table (bench.py, config 4)
        np.savez(out_path, out=allout.numpy().view(np.uint32), hash=np.uint64(h),
                 keysum=np.int64(sum(int(t.to(torch.int64).sum()) for t in dkeys.values())), names=np.array(sorted(dkeys)))
    else:
        assert allout is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("total", [37, 64])  # ragged and even shards
def test_world_size_2_gloo_matches_single_process(tmp_path, total):
    out_path = str(tmp_path / "w2.npz")
    mp.spawn(_worker, args=(2, _free_port(), total, out_path), nprocs=2, join=True)
    got = np.load(out_path)
    # the single-process run of the same job
    P = mk.derive_params(mk.PARAMSETS["STD100_MKNTRU_2"], mk.MKNTRU)
    rng = np.random.default_rng(11)
    keys = {"bk": rng.integers(0, P.Q, 5000, dtype=np.uint32), "pkey": rng.integers(0, P.Q, 300, dtype=np.uint32),
            "ksk": rng.integers(0, P.qKS, 777, dtype=np.uint32), "ctnand": rng.integers(0, P.q, P.k * P.n, dtype=np.uint32)}
    dkeys = shard.broadcast_keys(keys, "cpu", None)
    assert sorted(dkeys) == list(got["names"]) == ["bk", "ctnand", "ksk", "pkey"]  # absent blobs (kskA, kskB) are skipped
    assert int(got["keysum"]) == sum(int(t.to(torch.int64).sum()) for t in dkeys.values())
    one = _fake_gate(P, dkeys, shard.gate_inputs_torch(P, 4, 0, total, "cpu")).numpy().view(np.uint32)
    assert got["out"].shape == one.shape and (got["out"] == one).all()  # gathered in gate order
    assert int(got["hash"]) == shard.xor_hash_np(one, 0)  # digest independent of the number of ranks
