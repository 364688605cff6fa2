"""E-step of the unigram trainer on B200 (config 5, second half).

    python tools/estep_bench.py [--sentences 2000000] [--steps 5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P tools/estep_bench.py --gpus N

(word, freq) pairs = the whitespace-split words of the config-2 corpus with their counts
(what SplitSentencesByWhitespace hands to RunEStep), vocabulary = the pieces of the
unigram-32k fixture as a TrainerModel.  Words are sharded by bytes over the ranks; every
step ends with ONE all-reduce of [expected[V] | obj | ntokens] (NCCL).  Prints one JSON
line; the reference's E-step loop (oracle/_ref, all host threads) is timed next to it on
a bounded sample and the two results are compared (1e-5 relative on the log-likelihood).
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


def word_counts(n_sentences):
    from tools import synth
    text, off = synth.EnglishCorpus().generate(n_sentences)
    # sentences are packed back to back: put a space between them, then split on spaces
    b = bytearray(text.tobytes())
    raw = bytes(b)
    words = {}
    for i in range(len(off) - 1):
        for w in raw[off[i]:off[i + 1]].split(b" "):
            if w:
                words[w] = words.get(w, 0) + 1
    items = sorted(words.items(), key=lambda kv: (-kv[1], kv[0]))
    ws = [b"\xe2\x96\x81" + w for w, _ in items]
    freq = np.array([c for _, c in items], dtype=np.int64)
    offs = np.zeros(len(ws) + 1, dtype=np.int64)
    np.cumsum([len(w) for w in ws], out=offs[1:])
    return np.frombuffer(b"".join(ws), dtype=np.uint8).copy(), offs, freq, ws


RESULT_FD = 1


def main():
    global RESULT_FD
    sys.stdout.flush()
    RESULT_FD = os.dup(1)     # stdout carries the JSON line only; NCCL's banner (printf to fd 1) goes to stderr
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--sentences", type=int, default=2_000_000)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--cpu-words", type=int, default=200_000)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import spm_b200_pkg
    import protowire as pw
    from conftest import model_bytes
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)
    pkg = spm_b200_pkg.load()
    from spm_b200 import sharding
    ps = [(p, s) for p, s, t in pw.pieces_of(model_bytes("syn_uni32k")) if t == pw.NORMAL]
    pieces = [p for p, _ in ps]
    scores = np.array([s for _, s in ps], dtype=np.float32)
    text, off, freq, ws = word_counts(args.sentences)
    all_freq = float(freq.sum())
    lo, hi = sharding.estep_shard_ranges(off, freq, world)[rank]
    es = pkg.UnigramEStep(pieces, scores, device=local)
    d_text = torch.from_numpy(text[off[lo]:off[hi]]).to(dev)
    d_off = torch.from_numpy(off[lo:hi + 1] - off[lo]).to(dev)
    d_freq = torch.from_numpy(freq[lo:hi]).to(dev)
    V = len(pieces)
    packed = torch.zeros(V + 2, dtype=torch.float64, device=dev)     # expected | obj | ntokens
    stream = torch.cuda.current_stream(dev)

    def step():
        packed.zero_()
        es.run_device(d_text, d_off, d_freq, all_freq, packed[:V], packed[V:], stream=stream.cuda_stream)
        if world > 1:
            dist.all_reduce(packed, op=dist.ReduceOp.SUM)

    for _ in range(max(args.warmup, 3)):
        step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / args.steps
    if rank != 0:
        return
    res = packed.cpu().numpy()
    n_words = len(freq)
    line = {"metric": "unigram E-step words/s", "value": n_words / (ms * 1e-3), "unit": "words/s", "n_gpus": world,
            "steps": args.steps, "ms_per_step": ms, "scaling": "strong",
            "config": {"workload": "E-step: %d distinct words (%d B) of %d config-2 sentences, V=%d pieces "
                                   "(syn_uni32k as a TrainerModel)" % (n_words, int(off[-1]), args.sentences, V),
                       "exchange": "one all-reduce of %d float64 per step" % (V + 2) if world > 1 else "none (1 GPU)"},
            "obj": float(res[V]), "ntokens": int(round(res[V + 1])),
            "algorithmic_gb_per_s": (int(off[-1]) + 16 * n_words) / (ms * 1e-3) / 1e9}
    # reference on the host cores, bounded sample (most frequent words), threads as RunEStep uses them
    from oracle import ref
    if hasattr(ref.lib(), "ref_estep"):
        m = min(args.cpu_words, n_words)
        threads = max(1, len(os.sched_getaffinity(0)))
        parts = [None] * threads

        def work(k):
            sl = slice(k, m, threads)       # thread n takes words n, n + T, ... (:256)
            parts[k] = ref.estep(pieces, scores, ws[:m][sl], freq[:m][sl])
        t0 = time.perf_counter()
        th = [threading.Thread(target=work, args=(k,)) for k in range(threads)]
        for x in th:
            x.start()
        for x in th:
            x.join()
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": m / dt, "unit": "words/s", "cores": threads, "kind": "reference",
                                "sample": "first %d words, %.2f s" % (m, dt)}
        # parity on the sample: same words through the GPU
        e, o, nt, lz = es.run(text[:off[m]], off[:m + 1], freq[:m], want_logz=True)
        rexp = np.sum([p[0].astype(np.float64) for p in parts], axis=0)
        rnt = sum(p[3] for p in parts)
        big = rexp > 1e-2
        line["parity"] = {"ntokens_equal": bool(nt == rnt),
                          "max_rel_err_expected": float(np.max(np.abs(e[big] - rexp[big]) / rexp[big]))}
    sys.stdout.flush()
    os.write(RESULT_FD, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
