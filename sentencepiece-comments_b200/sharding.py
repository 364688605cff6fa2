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


def estep_shard_ranges(offsets, freq, world):
    """Contiguous (word, freq) ranges balanced by BYTES: the E-step visits every word once
    whatever its frequency, so the cost of a word is its length; cuts at
    k * total_bytes / world."""
    return shard_ranges(offsets, world)


def allreduce_estep(expected, obj, ntokens, group=None):
    """The only exchange step of the path: the serial merge of the per-thread
    accumulators in Trainer::RunEStep (src/unigram_model_trainer.cc:274-280) becomes ONE
    all-reduce(sum) of the packed vector [expected[V] | obj | ntokens] (float64).
    expected: numpy float64[V] (host) or a torch tensor; returns (expected, obj, ntokens)
    summed over the ranks, in the same container type."""
    import torch
    import torch.distributed as dist
    is_np = isinstance(expected, np.ndarray)
    e = torch.from_numpy(expected) if is_np else expected
    backend = dist.get_backend(group)
    dev = e.device if not is_np else torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else e.device
    packed = torch.empty(e.numel() + 2, dtype=torch.float64, device=dev)
    packed[:-2] = e.to(dev)
    packed[-2] = float(obj)
    packed[-1] = float(ntokens)
    dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=group)
    out = packed[:-2]
    res = out.cpu().numpy() if is_np else out
    return res, float(packed[-2].item()), int(round(float(packed[-1].item())))
