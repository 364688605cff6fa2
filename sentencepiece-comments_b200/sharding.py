"""Sentence sharding across the GPUs of one box.

Batch encode shards by independent units: each rank takes one contiguous
range of sentences, balanced by BYTES, and encodes it with its own model
replica.  There is no data-path collective; the only exchange is the
bookkeeping below (one all_gather of a per-rank id count) when a caller wants
global CSR offsets.  Works with any torch.distributed backend (nccl on the
GPUs, gloo in the CPU tests).
"""
import numpy as np


def shard_ranges(offsets, world):
    """offsets int64[n+1] -> [(lo, hi)] * world: contiguous sentence ranges whose byte
    counts are as equal as the sentence boundaries allow (cut at k * total / world)."""
    offsets = np.asarray(offsets, dtype=np.int64)
    n = len(offsets) - 1
    base, total = int(offsets[0]), int(offsets[-1] - offsets[0])
    cuts = [0]
    for k in range(1, world):
        target = base + (total * k) // world
        c = int(np.searchsorted(offsets, target, side="left"))
        cuts.append(min(max(c, cuts[-1]), n))
    cuts.append(n)
    return [(cuts[k], cuts[k + 1]) for k in range(world)]


def local_shard(text, offsets, rank, world):
    """This rank's (text view, rebased offsets, (lo, hi))."""
    lo, hi = shard_ranges(offsets, world)[rank]
    offsets = np.asarray(offsets, dtype=np.int64)
    off = offsets[lo:hi + 1]
    return text[off[0]:off[-1]], off - off[0], (lo, hi)


def global_id_offsets(local_id_offsets, group=None):
    """Per-rank CSR offsets (int64[n_r+1], starting at 0) -> this rank's offsets in the
    global ids array, plus the global id count.  One all_gather of one int64 per rank."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    mine = torch.tensor([int(local_id_offsets[-1])], dtype=torch.int64)
    backend = dist.get_backend(group)
    if backend == "nccl":
        mine = mine.cuda()
    allc = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(allc, mine, group=group)
    counts = [int(c.item()) for c in allc]
    base = sum(counts[:rank])
    return np.asarray(local_id_offsets, dtype=np.int64) + base, sum(counts)
