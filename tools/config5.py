#!/usr/bin/env python3
"""Config 5 of BASELINE.json as written: ONE corpus of 10^9 synthetic sentences, cut into
contiguous sentence ranges over the N GPUs of one box and STREAMED through
spm_cuda_encode_batch(SPM_CUDA_HOST) in blocks of 10 M sentences (~1 GB of text, ~0.9 GB of ids
per block) -- the regime the in-process loop `while (ReadLine) Encode` of
src/spm_encode_main.cc:183-193 turns into on a GPU box.

    python tools/config5.py [--blocks 100] [--block-sentences 10000000]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P tools/config5.py --gpus N

The corpus is "the 10 M base repeated" (SURVEY.md 8(d) config 5): block b of the corpus is the
config-2 generator's block (b mod K) with K distinct base blocks per rank held in pinned host
memory (generating 10^9 distinct sentences on the host would take longer than encoding them).
Rank r streams the contiguous block range shard_ranges() gives it (blocks are equal-sized, so
the byte-balanced cut is the block-count cut).  Double buffering: two worker threads per rank,
each with its own model handle (own device workspace, own copy streams) and its own pinned
input / output buffers, take alternate blocks -- the copy-in of block k+1 overlaps the copy-out
of block k on the full-duplex PCIe link.  No collective on the data path; torch.distributed only
carries the barrier and the MAX / SUM of the timings.

Next to it the host ceiling is MEASURED, not inferred: every rank runs pinned H2D and D2H copies
of 1 GiB at the same time as all other ranks (`pcie`), so `e2e_fraction_of_pcie_ceiling` says how
much of what the box's host memory / PCIe fabric can move the encode call actually moves.

Parity: the ids of the first block of rank 0 against the unmodified reference (oracle/_ref) on a
prefix, and every later repetition of a base block against the first one (total count, sampled
spans of ids and offsets).  One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def pcie_ceiling(ctx, nbytes=1 << 30, reps=4):
    """Per-rank pinned H2D / D2H GB/s, one direction and both at once, ALL ranks copying concurrently."""
    torch = ctx.torch
    h1 = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h2 = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d1 = torch.empty(nbytes, dtype=torch.uint8, device=ctx.dev)
    d2 = torch.empty(nbytes, dtype=torch.uint8, device=ctx.dev)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def both():
        with torch.cuda.stream(s1):
            d1.copy_(h1, non_blocking=True)
        with torch.cuda.stream(s2):
            h2.copy_(d2, non_blocking=True)

    out = {}
    for name, fn in (("h2d", lambda: d1.copy_(h1, non_blocking=True)), ("d2h", lambda: h2.copy_(d2, non_blocking=True)),
                     ("both", both)):
        fn()
        ctx.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        ctx.barrier()
        gbs = nbytes * reps / dt / 1e9
        mn = ctx.reduce([gbs], "MIN")[0]
        sm = ctx.reduce([gbs], "SUM")[0]
        out[name] = {"per_rank_min_gb_s": mn, "aggregate_gb_s": sm if name != "both" else 2 * sm}
    del h1, h2, d1, d2
    torch.cuda.empty_cache()
    return out


class Worker(threading.Thread):
    """One half of the double buffer: own handle, own pinned buffers, alternate blocks."""

    def __init__(self, ctx, mb, bases, n_ids_cap, blocks, options):
        super().__init__(daemon=True)
        torch = ctx.torch
        self.ctx, self.blocks = ctx, blocks
        self.sp = ctx.pkg.SentencePieceProcessor(device=ctx.local)
        self.sp.LoadFromSerializedProto(mb)
        for kv in options:
            k, _, v = kv.partition("=")
            self.sp.set_option(k, int(v))
        self.bases = bases            # [(pinned text, pinned offsets, n)]
        n = bases[0][2]
        self.h_ids = torch.empty(n_ids_cap, dtype=torch.int32).pin_memory()
        self.h_idoff = torch.empty(n + 1, dtype=torch.int64).pin_memory()
        self.stream = torch.cuda.Stream(device=ctx.dev)
        self.first = {}               # base index -> (total, sampled spans) of its first encoding
        self.mismatch = 0
        self.done_sentences = 0
        self.done_bytes = 0
        self.done_ids = 0
        self.err = None

    def check(self, k):
        n = self.bases[k][2]
        tot = int(self.h_idoff[n])
        spans = [(int(a), self.h_ids[int(a): int(a) + 4096].clone()) for a in np.linspace(0, max(0, tot - 4096), 7)]
        offs = self.h_idoff[:: max(1, n // 1000)].clone()
        if k not in self.first:
            self.first[k] = (tot, spans, offs)
            return
        t0, s0, o0 = self.first[k]
        if tot != t0 or not self.ctx.torch.equal(offs, o0) or any(
                not self.ctx.torch.equal(x[1], y[1]) for x, y in zip(spans, s0)):
            self.mismatch += 1

    def run(self):
        try:
            self.ctx.torch.cuda.set_device(self.ctx.local)     # (a new thread starts on device 0)
            with self.ctx.torch.cuda.stream(self.stream):
                for b in self.blocks:
                    k = b % len(self.bases)
                    text, off, n = self.bases[k]
                    self.sp.encode_pinned(text, off, self.h_ids, self.h_idoff, stream=self.stream.cuda_stream)
                    self.check(k)
                    self.done_sentences += n
                    self.done_bytes += int(off[n])
                    self.done_ids += int(self.h_idoff[n])
        except Exception as e:     # surfaces in the main thread
            self.err = e


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--blocks", type=int, default=100, help="blocks of the whole corpus (100 x 10 M = 10^9 sentences)")
    ap.add_argument("--block-sentences", type=int, default=10_000_000)
    ap.add_argument("--bases", type=int, default=2, help="distinct base blocks per rank kept in pinned memory")
    ap.add_argument("--model", default="syn_uni32k")
    ap.add_argument("--threads", type=int, default=2, help="streams per rank (2 = double buffer, 1 = serial calls)")
    ap.add_argument("--parity-sample", type=int, default=250_000)
    ap.add_argument("--option", action="append", default=[])
    ap.add_argument("--no-pcie", action="store_true")
    args = ap.parse_args()
    sys.stdout.flush()
    result = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ctx = bench.Ctx()
    torch = ctx.torch
    world, rank = ctx.world, ctx.rank
    from spm_b200 import sharding
    n = args.block_sentences
    # the corpus: `blocks` equal blocks; the byte-balanced contiguous cut over ranks (sharding.shard_ranges
    # on the block boundaries: equal blocks, so it is the block-count cut)
    block_off = np.arange(args.blocks + 1, dtype=np.int64) * n * 100
    lo, hi = sharding.shard_ranges(block_off, world)[rank]
    mb = bench.model_bytes(args.model)
    pcie = None if args.no_pcie else pcie_ceiling(ctx)
    bases = []
    for k in range(args.bases):
        text, off = bench.make_corpus("syn_en", n, (rank * args.bases + k) * n)
        bases.append((torch.from_numpy(text).pin_memory(), torch.from_numpy(off).pin_memory(), n))
    cap = max(int(b[1][n]) for b in bases) // 3 + 64       # ids <= bytes / 3 for this text (checked by the call)
    my_blocks = list(range(lo, hi))
    workers = [Worker(ctx, mb, bases, cap, my_blocks[t::args.threads], args.option) for t in range(args.threads)]
    # warm-up: every worker encodes every base once (workspace allocation, first-touch of the pinned outputs)
    for w in workers:
        with torch.cuda.stream(w.stream):
            for k in range(len(bases)):
                w.sp.encode_pinned(bases[k][0], bases[k][1], w.h_ids, w.h_idoff, stream=w.stream.cuda_stream)
                w.check(k)
    # parity of a sampled shard: rank 0, first block, prefix, against the unmodified reference
    parity = None
    if rank == 0 and args.parity_sample > 0:
        chk, kind = bench.ref_processor(mb)
        m = min(n, args.parity_sample)
        text, off, _ = bases[0]
        tnp, onp = text.numpy(), off.numpy()
        dt, wi, wo = bench.ref_encode(chk, kind, tnp[: onp[m]], onp[: m + 1], bench.host_threads() if kind == "reference" else 1)
        w0 = workers[0]
        with torch.cuda.stream(w0.stream):
            w0.sp.encode_pinned(bases[0][0], bases[0][1], w0.h_ids, w0.h_idoff, stream=w0.stream.cuda_stream)
        go = w0.h_idoff[: m + 1].numpy().copy()
        gi = w0.h_ids[: int(go[m])].numpy().copy()
        bad = int((np.diff(go) != np.diff(wo)).sum())
        if bad == 0 and not np.array_equal(gi, wi):
            bad = int(len(np.unique(np.searchsorted(wo, np.flatnonzero(gi != wi), side="right") - 1)))
        parity = {"sentences_compared": int(m), "sentences_mismatching": bad, "ids_compared": int(len(wi)), "checker": kind,
                  "reference_sentences_per_s": m / dt, "reference_threads": bench.host_threads()}
    ctx.barrier()
    t0 = time.perf_counter()
    for w in workers:
        w.start()
    for w in workers:
        w.join()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    ctx.barrier()
    for w in workers:
        if w.err:
            raise w.err
    wall_all = ctx.reduce([wall], "MAX")[0]
    sents, byts, ids, mism = ctx.reduce([float(sum(w.done_sentences for w in workers)),
                                         float(sum(w.done_bytes for w in workers)),
                                         float(sum(w.done_ids for w in workers)),
                                         float(sum(w.mismatch for w in workers))], "SUM")
    if rank == 0:
        moved = byts + 8 * sents + 4 * ids + 8 * sents       # H2D text + offsets, D2H ids + offsets
        line = {"metric": "config 5: streamed batch EncodeAsIds sentences/s (wall clock, pinned host buffers in and out)",
                "value": sents / wall_all, "unit": "sentences/s", "n_gpus": world, "wall_s": wall_all,
                "sentences": int(sents), "input_gb": byts / 1e9, "ids": int(ids),
                "input_gb_per_s": byts / wall_all / 1e9, "pcie_gb_per_s_moved": moved / wall_all / 1e9,
                "config": {"workload": "config5: %s, %d blocks x %d synthetic sentences, contiguous block ranges over %d GPU(s), "
                                       "%d stream(s) per GPU" % (args.model, args.blocks, n, world, args.threads),
                           "corpus": "%d base block(s) per rank repeated (SURVEY 8(d) config 5: 'the 10 M base repeated')" % args.bases,
                           "scaling": "strong (one corpus)"},
                "repeat_check": {"blocks_differing_from_their_first_encoding": int(mism)},
                "parity": parity, "pcie": pcie, "numa_node_rank0": ctx.numa}
        if pcie:
            ceil = pcie["both"]["aggregate_gb_s"]
            line["e2e_fraction_of_pcie_ceiling"] = (moved / wall_all / 1e9) / ceil if ceil > 0 else None
        print(json.dumps(line), file=result, flush=True)
    bad = (parity or {}).get("sentences_mismatching", 0) or mism
    if world > 1:
        ctx.dist.destroy_process_group()
    if bad:
        print("config5.py: PARITY FAILURE", file=sys.stderr)
        sys.exit(3)


if __name__ == "__main__":
    main()
