"""Multi-GPU host logic: gates shard across ranks, keys are replicated, results are gathered.

One process per GPU.  Every rank holds a replica of all parties' keys (rank 0 broadcasts the flattened blobs once; NCCL over
NVLink on the GPU box, gloo in the CPU tests), bootstraps the contiguous gate range `shard_range(total, world, rank)` and
the switched ciphertexts are gathered to rank 0.  There is no collective inside the data path: one gate is a sequential
recurrence over all parties' keys (mk-acc-xzw.cpp:253-275 of the reference sums over every party in every step), so only whole
gates are distributed.

Synthetic gate inputs are COUNTER-BASED: word w of input stream s of gate g is a pure function of (seed, g, s, w), so any
rank generates exactly its own shard, and any single gate can be regenerated on the CPU for the oracle check, whatever the
number of GPUs (BASELINE.json configs[3]).  The generator exists twice -- numpy (host) and torch (device) -- and the tests
prove the two agree.

Nothing here computes on ciphertexts; it only moves and indexes them.
"""
from __future__ import annotations

import numpy as np

_M64 = (1 << 64) - 1
_C_GATE, _C_STREAM, _C_WORD = 0x9E3779B97F4A7C15, 0xC2B2AE3D27D4EB4F, 0x165667B19E3779F9
_MIX1, _MIX2 = 0xBF58476D1CE4E5B9, 0x94D049BB133111EB


def shard_range(total: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous [lo, hi) of rank `rank`: sizes differ by at most one, earlier ranks take the remainder."""
    if world < 1 or not 0 <= rank < world or total < 0:
        raise ValueError("bad shard arguments")
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


# ---------------------------------------------------------------- counter-based inputs
def _splitmix_np(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64, copy=True)
    x ^= x >> np.uint64(30)
    x *= np.uint64(_MIX1)
    x ^= x >> np.uint64(27)
    x *= np.uint64(_MIX2)
    x ^= x >> np.uint64(31)
    return x


def counter_words_np(seed: int, gates: np.ndarray, stream: int, words: int, modulus: int) -> np.ndarray:
    """[len(gates), words] uint32 residues in [0, modulus): value = ((h >> 32) * modulus) >> 32 with h = splitmix64 of the
    counter.  `gates` are global gate indices."""
    with np.errstate(over="ignore"):
        g = np.asarray(gates, dtype=np.uint64)[:, None] * np.uint64(_C_GATE)
        w = np.arange(words, dtype=np.uint64)[None, :] * np.uint64(_C_WORD)
        base = np.uint64((seed * 0xD6E8FEB86659FD93 + stream * _C_STREAM) & _M64)
        h = _splitmix_np((g + w + base) ^ np.uint64(seed & _M64))
    return (((h >> np.uint64(32)) * np.uint64(modulus)) >> np.uint64(32)).astype(np.uint32)


def _s64(x: int) -> int:
    x &= _M64
    return x - (1 << 64) if x >= (1 << 63) else x


def counter_words_torch(seed: int, lo: int, hi: int, stream: int, words: int, modulus: int, device):
    """The same values as counter_words_np(seed, arange(lo, hi), ...), computed on `device` as an int32 tensor (bit pattern
    of the uint32 residue).  int64 arithmetic wraps like uint64; logical shifts are arithmetic shifts plus a mask."""
    import torch

    def lsr(x, s):
        return (x >> s) & ((1 << (64 - s)) - 1)

    g = torch.arange(lo, hi, dtype=torch.int64, device=device)[:, None] * _s64(_C_GATE)
    w = torch.arange(words, dtype=torch.int64, device=device)[None, :] * _s64(_C_WORD)
    base = _s64(seed * 0xD6E8FEB86659FD93 + stream * _C_STREAM)
    x = (g + w + base) ^ _s64(seed)
    x = x ^ lsr(x, 30)
    x = x * _s64(_MIX1)
    x = x ^ lsr(x, 27)
    x = x * _s64(_MIX2)
    x = x ^ lsr(x, 31)
    v = (lsr(x, 32) * modulus) >> 32  # < 2^32 * modulus < 2^63: no wrap
    return v.to(torch.int32) if modulus <= (1 << 31) else (v & 0xFFFFFFFF).to(torch.int32)


def gate_inputs_np(P, seed: int, gates) -> dict:
    """Inputs of the gates with global indices `gates`: ct1, ct2 [G][k][n] mod q and (MKLWE) b1, b2 [G] mod q."""
    gates = np.asarray(gates, dtype=np.int64)
    kn = P.k * P.n
    out = {"ct1": counter_words_np(seed, gates, 0, kn, P.q).reshape(-1, P.k, P.n),
           "ct2": counter_words_np(seed, gates, 1, kn, P.q).reshape(-1, P.k, P.n)}
    if P.scheme == 1:
        out["b1"] = counter_words_np(seed, gates, 2, 1, P.q).reshape(-1)
        out["b2"] = counter_words_np(seed, gates, 3, 1, P.q).reshape(-1)
    return out


def gate_inputs_torch(P, seed: int, lo: int, hi: int, device) -> dict:
    kn = P.k * P.n
    out = {"ct1": counter_words_torch(seed, lo, hi, 0, kn, P.q, device).view(-1, P.k, P.n),
           "ct2": counter_words_torch(seed, lo, hi, 1, kn, P.q, device).view(-1, P.k, P.n)}
    if P.scheme == 1:
        out["b1"] = counter_words_torch(seed, lo, hi, 2, 1, P.q, device).view(-1)
        out["b2"] = counter_words_torch(seed, lo, hi, 3, 1, P.q, device).view(-1)
    return out


# ---------------------------------------------------------------- key replication and result gather
KEY_NAMES = ("bk", "pkey", "ksk", "kskA", "kskB", "ctnand")


def broadcast_keys(keys: dict | None, device, dist=None, src: int = 0) -> dict:
    """Replicate the flattened key blobs of rank `src` on every rank as int32 tensors on `device`.

    keys: {name: uint32 numpy array} on rank `src` (None elsewhere).  Returns {name: tensor}.  With dist = None (one
    process) this is a plain upload.  The manifest (which blobs, how many words) travels first so receivers can allocate."""
    import torch

    rank = dist.get_rank() if dist is not None else 0
    manifest = torch.zeros(len(KEY_NAMES), dtype=torch.int64, device=device)
    if rank == src:
        for i, nme in enumerate(KEY_NAMES):
            if keys.get(nme) is not None:
                manifest[i] = int(np.asarray(keys[nme]).size)
    if dist is not None:
        dist.broadcast(manifest, src)
    out = {}
    for nme, words in zip(KEY_NAMES, manifest.tolist()):
        if words == 0:
            continue
        if rank == src:
            t = torch.from_numpy(np.ascontiguousarray(keys[nme]).reshape(-1).view(np.int32)).to(device)
        else:
            t = torch.empty(int(words), dtype=torch.int32, device=device)
        if dist is not None:
            dist.broadcast(t, src)
        out[nme] = t
    return out


class _DevMem:
    """A span of raw device memory as something torch.as_tensor understands (CUDA array interface, int32 words)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<i4", "data": (ptr, False), "version": 2}


def replicate_key_images(ctx, device, dist=None, src: int = 0) -> int:
    """Replicate the RESIDENT key images of rank `src`'s context (mkfhe_key_image: the keys as the kernels read them --
    Montgomery form, register order, u16 key-switch rows) into every other rank's context: one broadcast per image straight
    into the destination context's own HBM buffers (NCCL over NVLink / NVSwitch on the GPU box), no staging copy, no
    conversion on the receivers.  Rank `src` must have loaded its keys already.  Returns the bytes replicated per rank."""
    import torch

    total = 0
    images = ctx.KeyImages()
    for which, ptr, nbytes in images:
        assert nbytes % 4 == 0
        total += nbytes
        if dist is not None:
            t = torch.as_tensor(_DevMem(ptr, nbytes), device=device)
            dist.broadcast(t, src)
    if dist is not None and dist.get_rank() != src:
        torch.cuda.synchronize(device)
        ctx.CommitKeyImages([w for w, _, _ in images])
    return total


def gather_results(local, total: int, dist=None, dst: int = 0, out=None):
    """Rows of every rank's `local` tensor ([hi-lo, ...]) concatenated in gate order on rank `dst` (None elsewhere).
    Shards may be ragged (sizes differ by one): rows are padded to the largest shard for the collective.  A gather, not an
    all-gather: only `dst` holds the whole table (2^20 gates at k = 4 are 9.4 GB).  `out` (rank dst; optional): a tensor of
    world * largest-shard rows to receive into, reused across calls -- with even shards the result is a view of it and the
    call allocates and copies nothing."""
    import torch

    if dist is None:
        return local
    world, rank = dist.get_world_size(), dist.get_rank()
    sizes = [shard_range(total, world, r) for r in range(world)]
    most = max(hi - lo for lo, hi in sizes)
    pad = local
    if local.shape[0] < most:
        pad = torch.cat([local, local.new_zeros((most - local.shape[0],) + tuple(local.shape[1:]))])
    bufs = None
    if rank == dst:
        if out is None or out.shape[0] != world * most or out.dtype != local.dtype:
            out = torch.empty((world * most,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        bufs = list(out.chunk(world))
    dist.gather(pad.contiguous(), bufs, dst=dst)
    if rank != dst:
        return None
    if all(hi - lo == most for lo, hi in sizes):
        return out
    return torch.cat([b[: hi - lo] for b, (lo, hi) in zip(bufs, sizes)])


def gather_buffer(local, total: int, dist=None, dst: int = 0):
    """The reusable receive buffer gather_results(..., out=) wants on rank `dst` (None elsewhere / single process)."""
    import torch

    if dist is None or dist.get_rank() != dst:
        return None
    world = dist.get_world_size()
    most = max(hi - lo for lo, hi in (shard_range(total, world, r) for r in range(world)))
    return torch.empty((world * most,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)


def xor_hash_np(rows: np.ndarray, first_gate: int = 0) -> int:
    """Order- and sharding-independent digest of gate outputs: XOR over gates of splitmix64(gate index, words of the row).
    rows: [G, ...] uint32; row r belongs to global gate first_gate + r."""
    rows = np.ascontiguousarray(rows, dtype=np.uint32).reshape(len(rows), -1)
    with np.errstate(over="ignore"):
        acc = (np.arange(first_gate, first_gate + len(rows), dtype=np.uint64) + np.uint64(1)) * np.uint64(_C_GATE)
        for w in range(rows.shape[1]):  # sequential over words, vectorised over gates
            acc = _splitmix_np(acc ^ rows[:, w].astype(np.uint64)) + np.uint64(w)
    return int(np.bitwise_xor.reduce(acc)) if len(rows) else 0


def xor_hash_torch(rows, first_gate: int = 0, block: int = 1 << 16) -> int:
    """xor_hash_np on a torch tensor (int32 storage of the uint32 words), computed where the tensor lives -- the digest of a
    9 GB result table is taken on the GPU that gathered it.  Same value as xor_hash_np for the same rows."""
    import torch

    def lsr(x, s):
        return (x >> s) & ((1 << (64 - s)) - 1)

    def mix(x):
        x = x ^ lsr(x, 30)
        x = x * _s64(_MIX1)
        x = x ^ lsr(x, 27)
        x = x * _s64(_MIX2)
        return x ^ lsr(x, 31)

    rows = rows.reshape(rows.shape[0], -1)
    total = 0
    for a in range(0, rows.shape[0], block):
        r = rows[a:a + block]
        acc = (torch.arange(first_gate + a, first_gate + a + r.shape[0], dtype=torch.int64, device=rows.device) + 1) * _s64(_C_GATE)
        wide = r.to(torch.int64) & 0xFFFFFFFF
        for w in range(wide.shape[1]):
            acc = mix(acc ^ wide[:, w]) + w
        while acc.shape[0] > 1:  # XOR-reduce by halving
            half = acc.shape[0] // 2
            rest = acc[2 * half:]
            acc = torch.cat([acc[:half] ^ acc[half:2 * half], rest])
        total ^= int(acc[0].item()) & _M64 if acc.shape[0] else 0
    return total
