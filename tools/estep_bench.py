"""E-step of the unigram trainer on B200 (config 5, second half).

    python tools/estep_bench.py [--sentences 10000000] [--steps 5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P tools/estep_bench.py --gpus N

(word, freq) pairs = the whitespace-split words of the config-2 corpus with their counts
(what SplitSentencesByWhitespace hands to RunEStep, src/trainer_interface.cc:465-477),
against two vocabularies, each as a TrainerModel:
  final  the NORMAL pieces of the unigram-32k fixture (the last EM rounds);
  seed   >= 500,000 seed pieces made the way MakeSeedSentencePieces scores them (substrings of
         the words, freq x length, log-probabilities; src/unigram_model_trainer.cc:124-225) --
         the first EM rounds, where the lattices are dense and the step is long.
Words are sharded by bytes over the ranks; every step ends with ONE all-reduce of
[expected[V] | obj | ntokens] (NCCL, enqueued on the compute stream, no host sync between the
kernel and the collective).  `record()` is what bench.py puts under configs.estep; run as a
script it prints the same record as one JSON line.  The reference's E-step loop (oracle/_ref,
all host threads) is timed on a bounded sample and compared: per-word logZ and the float64
re-accumulation of the reference's own per-word posteriors at 1e-5 relative; the
delta to the reference's FLOAT accumulator is reported separately.
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
sys.path.insert(0, os.path.join(ROOT, "tests"))

WS = b"\xe2\x96\x81"


def word_counts(n_sentences):
    """((word bytes with the U+2581 prefix) list, freq int64[]) by descending frequency."""
    from tools import synth
    words, freq = synth.EnglishCorpus().word_counts(n_sentences)
    return [WS + w for w in words], np.asarray(freq, dtype=np.int64)


def pack(words):
    off = np.zeros(len(words) + 1, dtype=np.int64)
    np.cumsum([len(w) for w in words], out=off[1:])
    return np.frombuffer(b"".join(words), dtype=np.uint8).copy(), off


def seed_pieces(words, freq, n_seed=500_000, top_words=150_000, max_len=16):
    """Seed vocabulary in the manner of MakeSeedSentencePieces: every substring (1..16 characters,
    U+2581 only in front) of the most frequent words, scored freq x length, the best n_seed kept
    next to all single characters, scores = log probabilities."""
    m = min(top_words, len(words))
    # one byte per character: U+2581 -> 0x01, words separated by 0x00
    flat = b"\0".join(b"\x01" + w[3:] for w in words[:m]) + b"\0"
    a = np.frombuffer(flat, dtype=np.uint8)
    wid = np.cumsum(a == 0) - (a == 0)                  # word index of every byte
    f = np.asarray(freq[:m], dtype=np.float64)
    per_len = []                                          # (L, unique rows uint8[k, L], counts f64[k])
    for L in range(1, max_len + 1):
        if len(a) < L:
            break
        win = np.lib.stride_tricks.sliding_window_view(a, L)
        ok = (win != 0).all(axis=1)
        if L > 1:
            ok &= (win[:, 1:] != 1).all(axis=1)
        rows = np.ascontiguousarray(win[ok])
        if rows.size == 0:
            continue
        keys = rows.view(np.dtype((np.void, L))).ravel()
        uniq, inv = np.unique(keys, return_inverse=True)
        cnt = np.bincount(inv, weights=f[wid[: len(win)][ok]])
        per_len.append((L, uniq.view(np.uint8).reshape(-1, L), cnt))
    # all single characters, then the best multi-character substrings by freq x length (count >= 2)
    n_chars = len(per_len[0][2])
    score = np.concatenate([np.where(c >= 2, c * L, -1.0) for L, _, c in per_len[1:]])
    take = min(max(0, n_seed - n_chars), int((score > 0).sum()))
    chosen = np.argpartition(-score, take - 1)[:take] if take > 0 else np.zeros(0, dtype=np.int64)
    bounds = np.cumsum([0] + [len(c) for _, _, c in per_len[1:]])
    keys, counts = [bytes(r) for r in per_len[0][1]], list(per_len[0][2])
    which = np.searchsorted(bounds, chosen, side="right") - 1
    for g, j in zip(which.tolist(), (chosen - bounds[which]).tolist()):
        _, rows, c = per_len[1 + g]
        keys.append(rows[j].tobytes())
        counts.append(float(c[j]))
    counts = np.asarray(counts, dtype=np.float64)
    order = np.lexsort((np.arange(len(keys)), -counts))
    pieces = [keys[k].replace(b"\x01", WS) for k in order.tolist()]
    scores = np.log(counts[order] / counts.sum()).astype(np.float32)
    return pieces, scores


def final_pieces():
    import protowire as pw
    from conftest import model_bytes
    ps = [(p, s) for p, s, t in pw.pieces_of(model_bytes("syn_uni32k")) if t == pw.NORMAL]
    return [p for p, _ in ps], np.array([s for _, s in ps], dtype=np.float32)


def reference_sample(pieces, scores, ws, freq, m, threads):
    """The reference's loop on the first m words, `threads` host threads striding as RunEStep
    does (:256).  Returns (seconds, expected f32 (float accumulators summed), expected64, ntokens, logz)."""
    from oracle import ref
    parts = [None] * threads

    def work(k):
        sl = slice(k, m, threads)
        parts[k] = ref.estep64(pieces, scores, ws[:m][sl], freq[:m][sl])
    t0 = time.perf_counter()
    th = [threading.Thread(target=work, args=(k,)) for k in range(threads)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    dt = time.perf_counter() - t0
    exp32 = np.sum([p[0].astype(np.float64) for p in parts], axis=0)
    exp64 = np.sum([p[4] for p in parts], axis=0)
    nt = sum(p[3] for p in parts)
    logz = np.zeros(m, dtype=np.float32)
    for k in range(threads):
        logz[k:m:threads] = parts[k][1]
    return dt, exp32, exp64, nt, logz


def run_vocab(ctx, name, pieces, scores, ws, text, off, freq, steps, warmup, with_reference, cpu_words):
    torch, dist = ctx.torch, ctx.dist
    from spm_b200 import sharding
    world, rank, dev = ctx.world, ctx.rank, ctx.dev
    all_freq = float(freq.sum())
    lo, hi = sharding.estep_shard_ranges(off, freq, world)[rank]
    es = ctx.pkg.UnigramEStep(pieces, scores, device=ctx.local)
    d_text = torch.from_numpy(text[off[lo]:off[hi]]).to(dev)
    d_off = torch.from_numpy(off[lo:hi + 1] - off[lo]).to(dev)
    d_freq = torch.from_numpy(freq[lo:hi]).to(dev)
    V = len(pieces)
    packed = torch.zeros(V + 2, dtype=torch.float64, device=dev)     # expected | obj | ntokens
    stream = ctx.stream
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]

    def step(timed=False):
        packed.zero_()
        if timed:
            ev[0].record(stream)
        es.run_device(d_text, d_off, d_freq, all_freq, packed[:V], packed[V:], stream=stream.cuda_stream)
        if timed:
            ev[1].record(stream)
        if world > 1:
            dist.all_reduce(packed, op=dist.ReduceOp.SUM)     # on the current (compute) stream
        if timed:
            ev[2].record(stream)

    for _ in range(max(warmup, 3)):
        step()
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        step()
    e1.record(stream)
    ctx.barrier()
    ms = ctx.reduce([e0.elapsed_time(e1)], "MAX")[0] / steps
    step(timed=True)                      # one more step with events around its two halves
    ctx.barrier()
    k_ms, ar_ms = ctx.reduce([ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2])], "MAX")
    res = packed.cpu().numpy()
    n_words = len(freq)
    alg = int(off[-1]) + 16 * n_words + 8 * V
    rec = {"value": n_words / (ms * 1e-3), "unit": "words/s", "ms_per_step": ms,
           "workload": "E-step, %s vocabulary: V = %d pieces as a TrainerModel, %d distinct words (%d B)" % (
               name, V, n_words, int(off[-1])),
           "kernel_ms": k_ms, "allreduce_ms": ar_ms if world > 1 else 0.0,
           "allreduce_share_of_step": (ar_ms / (k_ms + ar_ms)) if world > 1 and k_ms + ar_ms > 0 else 0.0,
           "exchange": ("one all-reduce of %d float64 per step on the compute stream" % (V + 2)) if world > 1 else "none (1 GPU)",
           "obj": float(res[V]), "ntokens": int(round(res[V + 1])),
           "algorithmic_gb_per_s": alg / (ms * 1e-3) / 1e9}
    from bench import measured_peak
    peak, _ = measured_peak()
    rec["roofline"] = {"bound": "hbm", "achieved": alg / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0, "peak": peak,
                       "unit": "GB/s", "kernel": "k_estep*"}
    rec["roofline"]["frac"] = rec["roofline"]["achieved"] / peak
    if rank == 0 and with_reference:
        from oracle import ref
        if hasattr(ref.lib(), "ref_estep64"):
            threads = max(1, len(os.sched_getaffinity(0)))
            m = min(cpu_words, n_words)
            dt, r32, r64, rnt, rlogz = reference_sample(pieces, scores, ws, freq, m, threads)
            rec["cpu_baseline"] = {"value": m / dt, "unit": "words/s", "cores": threads, "kind": "reference",
                                   "sample": "first %d words, %.2f s (includes one TrainerModel build per thread)" % (m, dt)}
            e, o, nt, lz = es.run(text[:off[m]], off[:m + 1], freq[:m], want_logz=True)
            big = r64 > 1e-3 * max(1.0, float(r64.max()) * 1e-9)
            f32 = freq[:m].astype(np.float32)
            rec["parity"] = {
                "words_compared": int(m), "ntokens_equal": bool(nt == rnt),
                "max_rel_err_logz": float(np.max(np.abs(lz * f32 - rlogz) / np.maximum(np.abs(rlogz), 1e-30))),
                "max_rel_err_expected_vs_f64_reaccumulation": float(np.max(np.abs(e[big] - r64[big]) / r64[big])),
                "max_rel_err_expected_vs_reference_float_accumulator": float(np.max(np.abs(e[big] - r32[big]) / r32[big])),
                "tolerance": 1e-5}
            rec["parity"]["ok"] = bool(rec["parity"]["ntokens_equal"] and rec["parity"]["max_rel_err_logz"] <= 1e-5 and
                                       rec["parity"]["max_rel_err_expected_vs_f64_reaccumulation"] <= 1e-5)
    del es, d_text, d_off, d_freq, packed
    torch.cuda.empty_cache()
    return rec


def record(ctx, steps=5, warmup=3, sentences=10_000_000, with_reference=True, cpu_words=60_000, vocabs=("final", "seed")):
    ws, freq = word_counts(sentences)
    text, off = pack(ws)
    out = {"metric": "unigram E-step words/s", "n_gpus": ctx.world, "steps": steps, "scaling": "strong",
           "words_from": "%d config-2 sentences" % sentences}
    for name in vocabs:
        pieces, scores = final_pieces() if name == "final" else seed_pieces([w for w in ws], freq)
        out[name] = run_vocab(ctx, name, pieces, scores, ws, text, off, freq, steps, warmup, with_reference, cpu_words)
    # headline of the record: the seed-sized step (the long one)
    head = out.get("seed") or out.get("final")
    out["value"], out["unit"], out["ms_per_step"] = head["value"], head["unit"], head["ms_per_step"]
    return out


def main():
    sys.stdout.flush()
    result_fd = os.dup(1)     # stdout carries the JSON line only; NCCL's banner (printf to fd 1) goes to stderr
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--sentences", type=int, default=10_000_000)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--cpu-words", type=int, default=60_000)
    ap.add_argument("--vocabs", default="final,seed")
    ap.add_argument("--no-reference", action="store_true")
    args = ap.parse_args()
    import bench
    ctx = bench.Ctx()
    rec = record(ctx, steps=args.steps, warmup=args.warmup, sentences=args.sentences,
                 with_reference=not args.no_reference, cpu_words=args.cpu_words,
                 vocabs=tuple(v for v in args.vocabs.split(",") if v))
    if ctx.rank == 0:
        os.write(result_fd, (json.dumps(rec) + "\n").encode())
    if ctx.world > 1:
        ctx.dist.destroy_process_group()


if __name__ == "__main__":
    main()
