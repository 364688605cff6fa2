#!/usr/bin/env python3
"""Throughput of the steps either side of EncodeAsIds on one B200:

  --op pieces   spm_cuda_encode_batch_pieces   Encode(input, SentencePieceText*)
  --op decode   spm_cuda_decode_batch          Decode(ids, text)
  --op dense    spm_cuda_encode_batch + spm_cuda_csr_to_dense (the TF op's output layout)

One JSON line in bench.py's format: `value` device-resident (CUDA events on the launching
stream), `e2e` through the C ABI with pinned HOST buffers, `roofline` of the dominant stage,
`cpu_baseline` = the unmodified reference (oracle/_ref) on all host threads over a bounded
sample.  Algorithmic bytes per sentence: pieces: in + 8 + 20 ids + 8 + normalized + 8;
decode: 4 ids + 8 + text + 8; dense: as the id path + 4 max_len + 4.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def record(ctx, n=10_000_000, steps=3, parity_n=50_000, model="syn_uni32k", corpus="syn_en"):
    """Compact record for bench.py's `configs.pieces`: spm_cuda_encode_batch_pieces (Encode(input,
    SentencePieceText*), src/sentencepiece_processor.cc:488-573) device-resident on this rank's share of the
    config-2 corpus, timed with CUDA events; parity = serialized SentencePieceText of the first parity_n
    sentences against the unmodified reference's EncodeAsSerializedProto (rank 0)."""
    torch = ctx.torch
    pkg = ctx.pkg
    mb = bench.model_bytes(model)
    sp = pkg.SentencePieceProcessor(device=ctx.local)
    sp.LoadFromSerializedProto(mb)
    text, off = bench.make_corpus(corpus, n, ctx.rank * n)
    in_bytes = int(off[-1])
    dev = ctx.dev
    stream = torch.cuda.current_stream(dev)
    d_text, d_off = torch.from_numpy(text).to(dev), torch.from_numpy(off).to(dev)
    cap = in_bytes // 3 + 64
    arrs = [torch.empty(cap, dtype=torch.int32, device=dev) for _ in range(5)]
    d_idoff = torch.empty(n + 1, dtype=torch.int64, device=dev)
    d_norm = torch.empty(2 * in_bytes + 8 * n, dtype=torch.uint8, device=dev)
    d_normoff = torch.empty(n + 1, dtype=torch.int64, device=dev)

    def step():
        sp.encode_pieces_device(d_text, d_off, *arrs, d_idoff, d_norm, d_normoff, stream=stream.cuda_stream)
    for _ in range(3):
        step()
    stage_ms, launches = {}, 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.barrier()
    torch.cuda.synchronize()
    ev0.record(stream)
    for _ in range(steps):
        step()
        st = sp.last_stats()
        launches += st["launches"]
        for k, v in st["stage_ms"].items():
            stage_ms[k] = stage_ms.get(k, 0.0) + v
    ev1.record(stream)
    torch.cuda.synchronize()
    ms = ctx.reduce([ev0.elapsed_time(ev1)], "MAX")[0]
    n_ids = int(d_idoff[-1].item())
    rec = {"metric": "batch EncodeAsPieces / SentencePieceText sentences/s", "value": ctx.world * n * steps / (ms * 1e-3),
           "unit": "sentences/s", "ms_per_step": ms / steps, "steps": steps,
           "workload": "%s, %d %s sentences per GPU: ids + input / normalized spans + normalized text, device-resident"
                       % (model, n, corpus),
           "pieces_per_gpu": n_ids, "stage_ms_per_step": {k: round(v / steps, 4) for k, v in stage_ms.items()},
           "gpu_launches": launches}
    bad = 0
    if ctx.rank == 0 and parity_n > 0:
        chk, kind = bench.ref_processor(mb)
        if kind == "reference":
            m = min(n, parity_n)
            t0 = time.perf_counter()
            want, woff = chk.encode_proto_batch(text[: off[m]], off[: m + 1], threads=bench.host_threads())
            dt = time.perf_counter() - t0
            got = sp.encode_as_serialized_protos(text[: off[m]], off[: m + 1])
            bad = sum(1 for i in range(m) if got[i] != want[woff[i]: woff[i + 1]].tobytes())
            rec["parity"] = {"sentences_compared": m, "sentences_mismatching": bad, "checker": kind,
                             "what": "serialized SentencePieceText (pieces, ids, surfaces, begin / end)",
                             "reference_sentences_per_s": m / dt, "reference_threads": bench.host_threads()}
    del d_text, d_off, arrs, d_norm
    torch.cuda.empty_cache()
    bad_all = ctx.reduce([float(bad)], "SUM")[0]      # (every rank leaves the same way)
    if bad_all > 0:
        raise bench.ParityError("pieces: %d serialized protos of the compared prefix differ from the reference" % int(bad_all))
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--op", default="pieces", choices=["pieces", "decode", "dense"])
    ap.add_argument("--model", default="syn_uni32k")
    ap.add_argument("--corpus", default="syn_en", choices=["syn_en", "syn_dirty", "wagahai"])
    ap.add_argument("--sentences", type=int, default=2_000_000)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()

    import torch
    import spm_b200_pkg
    if not torch.cuda.is_available():
        raise SystemExit("pieces_bench.py: no CUDA device (this framework has no CPU path)")
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    pkg = spm_b200_pkg.load()
    mb = bench.model_bytes(args.model)
    sp = pkg.SentencePieceProcessor(device=0)
    sp.LoadFromSerializedProto(mb)
    n = args.sentences
    text, off = bench.make_corpus(args.corpus, n, 0)
    in_bytes = int(off[-1])
    stream = torch.cuda.current_stream(dev)
    d_text = torch.from_numpy(text).to(dev)
    d_off = torch.from_numpy(off).to(dev)
    d_ids = torch.empty(in_bytes + 64, dtype=torch.int32, device=dev)
    d_idoff = torch.empty(n + 1, dtype=torch.int64, device=dev)
    sp.encode_device(d_text, d_off, d_ids, d_idoff)
    n_ids = int(d_idoff[-1].item())
    h2d = d2h = 0

    if args.op == "pieces":
        cap = n_ids + 64
        arrs = [torch.empty(cap, dtype=torch.int32, device=dev) for _ in range(5)]
        ncap = 2 * in_bytes + 8 * n
        d_norm = torch.empty(ncap, dtype=torch.uint8, device=dev)
        d_normoff = torch.empty(n + 1, dtype=torch.int64, device=dev)

        def step():
            sp.encode_pieces_device(d_text, d_off, *arrs, d_idoff, d_norm, d_normoff, stream=stream.cuda_stream)
        step()
        n_norm = int(d_normoff[-1].item())
        alg = in_bytes + 8 * n + 20 * n_ids + 8 * n + n_norm + 8 * n
        h = {"text": torch.from_numpy(text).pin_memory(), "off": torch.from_numpy(off).pin_memory(),
             "arrs": [torch.empty(cap, dtype=torch.int32).pin_memory() for _ in range(5)],
             "idoff": torch.empty(n + 1, dtype=torch.int64).pin_memory(),
             "norm": torch.empty(n_norm + 64, dtype=torch.uint8).pin_memory(),
             "normoff": torch.empty(n + 1, dtype=torch.int64).pin_memory()}

        def step_host():
            pkg._check(pkg.lib().spm_cuda_encode_batch_pieces(
                sp._h, h["text"].data_ptr(), h["off"].data_ptr(), n, 0, *[a.data_ptr() for a in h["arrs"]], cap,
                h["idoff"].data_ptr(), h["norm"].data_ptr(), h["norm"].numel(), h["normoff"].data_ptr(), pkg.HOST,
                stream.cuda_stream))
        h2d, d2h = in_bytes + 8 * (n + 1), 20 * n_ids + n_norm + 16 * (n + 1)
        what = "Encode(input, SentencePieceText*): ids + input / normalized spans + normalized text"
    elif args.op == "decode":
        d_out = torch.empty(2 * in_bytes + 64, dtype=torch.uint8, device=dev)
        d_toff = torch.empty(n + 1, dtype=torch.int64, device=dev)

        def step():
            sp.decode_device(d_ids[:n_ids], d_idoff, d_out, d_toff, stream=stream.cuda_stream)
        step()
        n_text = int(d_toff[-1].item())
        alg = 4 * n_ids + 8 * n + n_text + 8 * n
        h = {"ids": d_ids[:n_ids].cpu().pin_memory(), "idoff": d_idoff.cpu().pin_memory(),
             "text": torch.empty(n_text + 64, dtype=torch.uint8).pin_memory(),
             "toff": torch.empty(n + 1, dtype=torch.int64).pin_memory()}

        def step_host():
            pkg._check(pkg.lib().spm_cuda_decode_batch(sp._h, h["ids"].data_ptr(), h["idoff"].data_ptr(), n,
                                                       h["text"].data_ptr(), h["text"].numel(), h["toff"].data_ptr(),
                                                       pkg.HOST, stream.cuda_stream))
        h2d, d2h = 4 * n_ids + 8 * (n + 1), n_text + 8 * (n + 1)
        what = "Decode(ids) of the batch's own ids"
    else:
        def step():
            sp.encode_device(d_text, d_off, d_ids, d_idoff, stream=stream.cuda_stream)
            return sp.csr_to_dense(d_ids, d_idoff, pad=0, stream=stream.cuda_stream)
        v, _ = step()
        alg = in_bytes + 8 * n + 4 * v.shape[1] * n + 4 * n
        step_host = None
        what = "EncodeAsIds + dense [batch, max_len=%d] / sequence_length layout" % v.shape[1]

    for _ in range(max(args.warmup, 3)):
        step()
    stage_ms = {}
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    launches = 0
    with bench.ClockSampler(0) as clk:
        ev0.record(stream)
        for _ in range(args.steps):
            step()
            st = sp.last_stats()
            launches += st["launches"]
            for k, v in st["stage_ms"].items():
                stage_ms[k] = stage_ms.get(k, 0.0) + v
        ev1.record(stream)
        torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    dom = max(stage_ms, key=stage_ms.get)
    peak, peak_src = bench.measured_peak()
    k_ms = stage_ms[dom] / args.steps
    kernel = {"pieces": {"encode": "k_gen_encode (exact tier with piece ends)", "normalize": "k_pc_normalize",
                         "compact": "k_scan_* + k_pc_emit + k_pc_copy_norm"},
              "decode": {"encode": "k_decode<count> + k_scan_* + k_decode<write>"},
              "dense": bench.STAGE_KERNEL}[args.op].get(dom, dom)
    line = {"metric": "batch %s sentences/s" % args.op, "value": n * args.steps / (ms * 1e-3), "unit": "sentences/s",
            "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8 text / int32 ids and spans / f32 scores", "data": "synthetic",
            "config": {"workload": "%s: %s, %d %s sentences" % (what, args.model, n, args.corpus),
                       "input_bytes": in_bytes, "ids": n_ids,
                       "l2": "inputs (%.2f GB) larger than L2; no flush" % (in_bytes / 1e9)},
            "input_gb_per_s": in_bytes * args.steps / (ms * 1e-3) / 1e9,
            "stage_ms_per_step": {k: round(v / args.steps, 4) for k, v in stage_ms.items()},
            "roofline": {"bound": "hbm", "achieved": alg / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0, "peak": peak,
                         "unit": "GB/s", "frac": (alg / (k_ms * 1e-3) / 1e9 / peak) if k_ms > 0 else 0.0,
                         "traffic": None, "kernel": kernel, "kernel_ms_per_step": k_ms,
                         "kernel_share_of_step": stage_ms[dom] / ms if ms > 0 else None,
                         "algorithmic_bytes_per_step": alg, "peak_source": peak_src},
            "clocks": clk.summary(), "gpu_launches": launches}
    if step_host is not None and not args.no_e2e:
        for _ in range(2):
            step_host()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(stream)
        for _ in range(args.steps):
            step_host()
        e1.record(stream)
        torch.cuda.synchronize()
        ems = e0.elapsed_time(e1)
        line["e2e"] = {"value": n * args.steps / (ems * 1e-3), "unit": "sentences/s", "h2d_bytes_per_step": h2d,
                       "d2h_bytes_per_step": d2h, "ms_per_step": ems / args.steps}
    if not args.no_cpu_baseline and args.op != "dense":
        chk, kind = bench.ref_processor(mb)
        if kind == "reference":
            threads = bench.host_threads()
            m = min(n, 60_000 * threads)
            t0 = time.perf_counter()
            if args.op == "pieces":
                chk.encode_pieces_count(text[: off[m]], off[: m + 1], threads=threads)
            else:
                hid = d_ids[:n_ids].cpu().numpy()
                hio = d_idoff.cpu().numpy()
                chk.decode_count(hid[: hio[m]], hio[: m + 1], threads=threads)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": m / dt, "unit": "sentences/s", "cores": threads, "kind": kind,
                                    "sample": "first %d sentences of the workload, %.1f s wall, in-process %s loop" % (
                                        m, dt, "Encode(SentencePieceText)" if args.op == "pieces" else "Decode(ids)")}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
