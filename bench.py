#!/usr/bin/env python3
"""bench.py -- batch EncodeAsIds throughput on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One "step" = one pass of spm_cuda_encode_batch over one batch of synthetic
sentences.  The headline line is config 2 of BASELINE.json (unigram-32k,
nmt_nfkc, 10 M English-like sentences of ~100 B per GPU; sentence-sharded, no
collective, weak scaling):
  `value`  the device-resident call (CUDA events on the launching stream);
  `e2e`    the same call with HOST (pinned) buffers, H2D/D2H inside the timed region;
  `parity` ids of the timed call compared with the unmodified reference
           (oracle/_ref) on a prefix of the benched batch -- a mismatch is a non-zero exit;
  `configs` the same record for config 3 (BPE-32k), config 4 (CJK lines, >= 1 GB),
           config 2 with dirty text, the unigram E-step (with its all-reduce at N > 1) and the
           pieces call (EncodeAsPieces / SentencePieceText, serialized protos compared with the
           reference's), measured in the same process after the headline run.
`--impl reference` times the unmodified reference compiled into oracle/_ref (all host
threads) on a bounded sample of the same workload.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "batch EncodeAsIds sentences/s"
UNIT = "sentences/s"
MODELS = os.path.join(ROOT, "tests", "golden", "models")
CSRC = os.path.join(ROOT, "sentencepiece-comments_b200", "csrc")
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "traffic.json")

# the extra records of the one JSON line: name -> (model, corpus, sentences per GPU, parity sample at N = 1)
EXTRA_CONFIGS = {
    "config3": ("syn_bpe32k", "syn_en", 10_000_000, 1_000_000),
    "config4": ("ref_ja_8k", "wagahai", 2_200_000, 60_000),
    "config2_dirty": ("syn_uni32k", "syn_dirty", 10_000_000, 1_000_000),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="syn_uni32k")
    ap.add_argument("--sentences", type=int, default=10_000_000, help="sentences per GPU per step")
    ap.add_argument("--corpus", default="syn_en", choices=["syn_en", "syn_dirty", "wagahai"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="sentences of the parity / cpu_baseline sample (0: auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the reference leg (no parity counters either)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--configs", default="config3,config4,config2_dirty,estep,pieces",
                    help="extra records measured after the headline run ('' or --no-configs: none)")
    ap.add_argument("--no-configs", action="store_true")
    ap.add_argument("--option", action="append", default=[], help="name=value for spm_cuda_model_set_option (A/B runs)")
    return ap.parse_args()


def model_bytes(name):
    path = name if os.path.exists(name) else os.path.join(MODELS, name + ".model")
    with open(path, "rb") as f:
        return f.read()


_CORPORA = {}


def make_corpus(kind, n, first):
    from tools import synth
    if kind not in _CORPORA:
        _CORPORA[kind] = (synth.EnglishCorpus() if kind == "syn_en" else
                          synth.EnglishCorpus(dirty=True) if kind == "syn_dirty" else synth.ResampledCorpus())
    return _CORPORA[kind].generate(n, first=first)


def workload_name(model, corpus, n):
    return {"syn_en": "config2: %s, %d synthetic English-like sentences (~100 B) per GPU",
            "syn_dirty": "config2-dirty: %s, %d synthetic sentences with NFKC/whitespace/malformed noise per GPU",
            "wagahai": "config4: %s, %d wagahai-style CJK lines per GPU"}[corpus] % (model, n)


def config_of(model, corpus, n, gpus):
    """`config` of the JSON line: a pure function of the arguments, so that the b200 arm and the
    reference arm print the very same object for the same command line."""
    return {"workload": workload_name(model, corpus, n), "sentences_per_gpu": n,
            "l2": "inputs (~%.1f GB per GPU) larger than L2; no flush" % (n * (100 if corpus != "wagahai" else 477) / 1e9),
            "parallelism": "sentence-sharded x%d, no collective" % gpus}


class ClockSampler:
    """SM clock / throttle reasons sampled DURING the timed region (NVML every few ms;
    nvidia-smi as the fallback when the NVML module is missing)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.index = index
        self.sm, self.mx, self.reasons = [], 0.0, set()
        self.stop = False
        self.t = None
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            uuid = None
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = index
            if vis:
                ent = vis.split(",")[index].strip()
                if ent.isdigit():
                    idx = int(ent)
                else:
                    uuid = ent
            self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid) if uuid else pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n = self.nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)))
        try:
            r = n.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        for bit, nm in ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"),
                        (0x4, "sw_power_cap")):
            if r & bit:
                self.reasons.add(nm)

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                              "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10)
        if out.returncode == 0 and out.stdout.strip():
            r = [x.strip() for x in out.stdout.strip().split("\n")[0].split(",")]
            self.sm.append(float(r[0]))
            self.mx = max(self.mx, float(r[1]))
            for k, nm in enumerate(self.NAMES):
                if len(r) > 3 + k and r[3 + k].lower().startswith("active"):
                    self.reasons.add(nm)

    def _run(self):
        while not self.stop:
            try:
                if self.nvml:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            time.sleep(0.004 if self.nvml else 0.05)

    def __enter__(self):
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=15)

    def summary(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx or None,
                "reasons": sorted(self.reasons), "samples": len(self.sm),
                "source": "nvml" if self.nvml else "nvidia-smi"}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def source_hash():
    """sha256 over the kernel sources: a dram-traffic figure captured with ncu is only quoted
    next to the kernels it was captured from."""
    h = hashlib.sha256()
    for fn in sorted(os.listdir(CSRC)):
        with open(os.path.join(CSRC, fn), "rb") as f:
            h.update(fn.encode() + b"\0" + f.read())
    return h.hexdigest()[:16]


def recorded_traffic(stage, model, corpus, n):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the stage's kernel, from the
    `ncu --set full` capture summarised by tools/ncu_traffic.py into profiles/traffic.json --
    None unless that capture was taken from these very sources on this very workload."""
    try:
        with open(TRAFFIC_FILE) as f:
            tab = json.load(f)
    except (OSError, ValueError):
        return None, None
    e = tab.get("%s|%s|%s|%d" % (stage, model, corpus, n))
    if not e or e.get("src_sha") != source_hash():
        return None, ("stale: " + e.get("source", "?")) if e else None
    return float(e["traffic"]), e.get("source")


def ref_processor(mb):
    """The CPU arm: the unmodified reference (oracle/_ref) when its library is
    present, else the CPU restatement (oracle port; single-threaded)."""
    from oracle import port, ref
    if ref.available():
        return ref.RefProcessor(mb), "reference"
    return port.OracleProcessor(mb), "port"


def ref_encode(chk, kind, text, off, threads):
    """(seconds, ids, id_offsets) of the reference's Encode(ids) loop over the sample."""
    t0 = time.perf_counter()
    if kind == "reference":
        ids, ido = chk.encode_batch(text, off, threads=threads)
    else:
        ids, ido = chk.encode_batch(text, off)
    return time.perf_counter() - t0, ids, ido


def run_reference(args):
    """`--impl reference`: the reference's own CPU Encode loop over a bounded
    sample of the same workload, all host threads; rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    mb = model_bytes(args.model)
    chk, kind = ref_processor(mb)
    threads = host_threads() if kind == "reference" else 1
    n = args.cpu_sample or min(args.sentences, 40_000 * threads)
    text, off = make_corpus(args.corpus, n, 0)
    for _ in range(args.warmup):
        ref_encode(chk, kind, text, off, threads)
    t = 0.0
    n_ids = 0
    for _ in range(args.steps):
        dt, ids, _ = ref_encode(chk, kind, text, off, threads)
        t += dt
        n_ids = len(ids)
    v = n * args.steps / t
    sample = "first %d sentences of the workload per step (%d B, %d ids), in-process Encode(ids) loop, no file I/O" % (
        n, int(off[-1]), n_ids)
    # `config` is byte-identical to the b200 arm's; the sample is described under cpu_baseline only
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32 ids, f32 scores",
            "data": "synthetic", "config": config_of(args.model, args.corpus, args.sentences, args.gpus),
            "input_gb_per_s": float(off[-1]) * args.steps / t / 1e9,
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), file=RESULT, flush=True)


def bind_to_gpu_numa_node(index):
    """Run this rank (and first-touch its pinned buffers) on the NUMA node its GPU hangs off:
    with one process per GPU the e2e leg is bound by host memory / PCIe root-complex traffic."""
    try:
        import torch
        p = torch.cuda.get_device_properties(index)
        dom, bus, devn = getattr(p, "pci_domain_id", 0), p.pci_bus_id, p.pci_device_id
        with open("/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bus, devn)) as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        cpus = set()
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            for part in f.read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return node
    except Exception:
        pass
    return None


STAGE_KERNEL = {"encode": "k_encode_fast / k_encode_warp / k_encode_bpe_fast (general tier kernels for models without a fast tier)",
                "normalize": "k_normalize_fast / k_normalize_long",
                "memo_collect": "k_memo_scan_words (+ k_memo_scan for the listed sentences)",
                "memo_words": "k_memo_words_*", "memo_emit": "k_memo_emit", "compact": "k_scan_* + k_compact"}


class Ctx:
    """Process-wide state of one rank."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (this framework has no CPU path)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
                os.environ["NCCL_DEBUG"] = "WARN"      # keep stdout to the one JSON line
            dist.init_process_group("nccl", device_id=self.dev)
        self.numa = bind_to_gpu_numa_node(self.local) if self.world > 1 else None
        import spm_b200_pkg
        self.pkg = spm_b200_pkg.load()
        self.stream = torch.cuda.current_stream(self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, vals, op):
        t = self.torch.tensor(vals, dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=getattr(self.dist.ReduceOp, op))
        return t.tolist()


class ParityError(Exception):
    pass


def measure_encode(ctx, model, corpus, n, steps, warmup, parity_n, options=(), want_e2e=True):
    """One encode workload on this rank's shard; returns the record (rank 0: complete)."""
    torch = ctx.torch
    dev, stream = ctx.dev, ctx.stream
    mb = model_bytes(model)
    sp = ctx.pkg.SentencePieceProcessor(device=ctx.local)
    sp.LoadFromSerializedProto(mb)
    for kv in options:
        k, _, v = kv.partition("=")
        sp.set_option(k, int(v))
    text, off = make_corpus(corpus, n, ctx.rank * n)   # this rank's shard (contiguous sentence range)
    in_bytes = int(off[-1])
    d_text = torch.from_numpy(text).to(dev)
    d_off = torch.from_numpy(off).to(dev)
    d_ids = torch.empty(in_bytes + 64, dtype=torch.int32, device=dev)   # <= one id per input byte
    d_idoff = torch.empty(n + 1, dtype=torch.int64, device=dev)

    def step_device():
        sp.encode_device(d_text, d_off, d_ids, d_idoff, stream=stream.cuda_stream)
        return sp.last_stats()

    st = None
    for _ in range(max(warmup, 3)):
        st = step_device()
    n_ids = st["ids"]
    alg_bytes = in_bytes + 8 * n + 4 * n_ids + 8 * n   # per sentence in + 8 + 4 ids + 8

    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    stage_ms = {}
    ctx.barrier()
    with ClockSampler(ctx.local) as clk:
        ev0.record(stream)
        for _ in range(steps):
            st = step_device()
            launches += st["launches"]
            for k, v in st["stage_ms"].items():
                stage_ms[k] = stage_ms.get(k, 0.0) + v
        ev1.record(stream)
        ctx.barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = clk.summary()
    # the dominant stage of the step: one kernel per call (two when the batch holds lines >= 4 KiB)
    dom = max(stage_ms, key=stage_ms.get)
    kern_ms, kern_launches = stage_ms[dom], steps

    # ---- e2e: same call, HOST (pinned) buffers, copies inside the timed region
    e2e = None
    host_equals_device = None
    if want_e2e:
        h_text = torch.from_numpy(text).pin_memory()
        h_off = torch.from_numpy(off).pin_memory()
        h_ids = torch.empty(n_ids + 64, dtype=torch.int32).pin_memory()
        h_idoff = torch.empty(n + 1, dtype=torch.int64).pin_memory()
        for _ in range(2):
            sp.encode_pinned(h_text, h_off, h_ids, h_idoff, stream=stream.cuda_stream)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ctx.barrier()
        e0.record(stream)
        for _ in range(steps):
            sp.encode_pinned(h_text, h_off, h_ids, h_idoff, stream=stream.cuda_stream)
        e1.record(stream)
        ctx.barrier()
        e2e_ms = e0.elapsed_time(e1)
        e2e = (e2e_ms, in_bytes + 8 * (n + 1), 4 * n_ids + 8 * (n + 1))
        # the HOST call (pipelined chunks) and the DEVICE call (one chunk) produced the same CSR, every id
        host_equals_device = bool(int(h_idoff[n]) == n_ids and
                                  torch.equal(h_idoff.to(dev), d_idoff) and
                                  torch.equal(h_ids[:n_ids].to(dev), d_ids[:n_ids]))
        del h_text, h_off, h_ids, h_idoff

    # ---- parity counters: the ids of the timed device call against the reference, sentence by sentence
    parity, cpu = None, None
    if ctx.rank == 0 and parity_n > 0:
        chk, kind = ref_processor(mb)
        threads = host_threads() if kind == "reference" else 1
        m = min(n, parity_n)
        dt, wi, wo = ref_encode(chk, kind, text[: off[m]], off[: m + 1], threads)
        go = d_idoff[: m + 1].cpu().numpy()
        gi = d_ids[: int(go[m])].cpu().numpy()
        same_counts = np.diff(go) == np.diff(wo)
        bad = int((~same_counts).sum())
        if bad == 0 and not np.array_equal(gi, wi):
            neq = np.flatnonzero(gi != wi)
            bad = int(len(np.unique(np.searchsorted(wo, neq, side="right") - 1)))
        parity = {"sentences_compared": int(m), "sentences_mismatching": bad, "ids_compared": int(len(wi)),
                  "checker": kind, "host_call_equals_device_call": host_equals_device}
        cpu = {"value": m / dt, "unit": UNIT, "cores": threads, "kind": kind,
               "sample": "first %d sentences of the workload (%d B, %d ids), %.1f s wall, in-process "
                         "Encode(ids) loop; its ids are the parity check" % (m, int(off[m]), len(wi), dt)}
        del chk

    ms_all, e2e_all = ctx.reduce([ms, e2e[0] if e2e else 0.0], "MAX")
    n_all, bytes_all, ids_all = ctx.reduce([float(n), float(in_bytes), float(n_ids)], "SUM")
    ok_all = ctx.reduce([0.0 if host_equals_device is False else 1.0], "MIN")[0]
    bad_all = ctx.reduce([float(parity["sentences_mismatching"]) if parity else 0.0], "MAX")[0]   # every rank learns it

    peak, peak_src = measured_peak()
    traffic, traffic_src = recorded_traffic(dom, model, corpus, n)
    k_avg_ms = kern_ms / max(kern_launches, 1)
    alg_per_launch = alg_bytes * steps / max(kern_launches, 1)
    achieved = alg_per_launch / (k_avg_ms * 1e-3) / 1e9 if k_avg_ms > 0 else 0.0
    rec = {"value": n_all * steps / (ms_all * 1e-3), "unit": UNIT, "ms_per_step": ms_all / steps,
           "config": config_of(model, corpus, n, ctx.world),
           "workload_detail": {"input_bytes_per_gpu": in_bytes, "ids_per_gpu": n_ids,
                               "numa_node_rank0": ctx.numa},
           "input_gb_per_s": bytes_all * steps / (ms_all * 1e-3) / 1e9,
           "ids_per_s": ids_all * steps / (ms_all * 1e-3),
           "general_tier_sentences_per_step": st["fallback_sentences"],
           "viterbi_tier_sentences_per_step": st["viterbi_tier_sentences"],
           "stage_ms_per_step": {k: round(v / steps, 4) for k, v in stage_ms.items()},
           "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                        "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                        "kernel": STAGE_KERNEL.get(dom, dom),
                        "kernel_ms_per_launch": k_avg_ms, "kernel_launches": kern_launches,
                        "kernel_share_of_step": kern_ms / ms if ms > 0 else None,
                        "algorithmic_bytes_per_launch": alg_per_launch,
                        "whole_call_frac": alg_bytes / (ms / steps * 1e-3) / 1e9 / peak, "peak_source": peak_src},
           "clocks": clocks, "gpu_launches": launches}
    if e2e:
        rec["e2e"] = {"value": n_all * steps / (e2e_all * 1e-3), "unit": UNIT,
                      "h2d_bytes_per_step": e2e[1], "d2h_bytes_per_step": e2e[2], "ms_per_step": e2e_all / steps}
    if parity:
        rec["parity"] = parity
        rec["cpu_baseline"] = cpu
    del d_text, d_off, d_ids, d_idoff, sp
    torch.cuda.empty_cache()
    if ok_all < 1.0:
        raise ParityError("%s: the HOST call and the DEVICE call disagree" % workload_name(model, corpus, n))
    if bad_all > 0:
        raise ParityError("%s: %d sentences of the compared prefix differ from the reference" % (
            workload_name(model, corpus, n), int(bad_all)))
    return rec


def compact_record(rec):
    """What an extra config contributes to the one JSON line."""
    out = {k: rec[k] for k in ("value", "unit", "ms_per_step", "input_gb_per_s", "stage_ms_per_step",
                               "general_tier_sentences_per_step", "viterbi_tier_sentences_per_step",
                               "gpu_launches") if k in rec}
    out["workload"] = rec["config"]["workload"]
    out["input_bytes_per_gpu"] = rec["workload_detail"]["input_bytes_per_gpu"]
    if "e2e" in rec:
        out["e2e"] = rec["e2e"]
    r = rec["roofline"]
    out["roofline"] = {k: r[k] for k in ("frac", "achieved", "peak", "unit", "kernel", "kernel_ms_per_launch",
                                         "traffic", "whole_call_frac")}
    for k in ("parity", "cpu_baseline"):
        if k in rec:
            out[k] = rec[k]
    out["clocks"] = {k: rec["clocks"][k] for k in ("sm_mhz", "reasons")}
    return out


def run_b200(args):
    ctx = Ctx()
    world, rank = ctx.world, ctx.rank
    steps, warmup = args.steps, max(args.warmup, 3)
    auto_sample = 2_000_000 if world == 1 else 250_000
    parity_n = 0 if args.no_cpu_baseline else (args.cpu_sample or auto_sample)
    failed = None
    try:
        rec = measure_encode(ctx, args.model, args.corpus, args.sentences, steps, warmup, parity_n,
                             options=args.option, want_e2e=not args.no_e2e)
    except ParityError as e:
        failed = str(e)
        rec = None
    extras = {}
    names = [] if (args.no_configs or failed) else [c for c in args.configs.split(",") if c]
    for name in names:
        t0 = time.perf_counter()
        try:
            if name == "estep":
                from tools import estep_bench
                extras[name] = estep_bench.record(ctx, steps=max(3, min(steps, 5)), with_reference=parity_n > 0)
            elif name == "pieces":
                from tools import pieces_bench
                pn = 10_000_000 if args.sentences >= 10_000_000 else max(1000, args.sentences)
                extras[name] = pieces_bench.record(ctx, n=pn, steps=3, parity_n=50_000 if (parity_n > 0 and world == 1) else
                                                   (12_500 if parity_n > 0 else 0))
            elif name in EXTRA_CONFIGS:
                model, corpus, n, pn = EXTRA_CONFIGS[name]
                if args.sentences < 10_000_000:           # a scaled-down run scales the extras with it
                    n = max(1000, int(n * args.sentences / 10_000_000))
                pn = 0 if parity_n == 0 else (pn if world == 1 else pn // 4)
                extras[name] = compact_record(measure_encode(ctx, model, corpus, n, max(3, min(steps, 5)), 3, pn,
                                                             options=args.option, want_e2e=not args.no_e2e))
            else:
                extras[name] = {"error": "unknown config"}
        except ParityError as e:
            failed = str(e)
            extras[name] = {"error": failed}
            break
        except Exception as e:   # an extra record never takes the headline down with it
            extras[name] = {"error": "%s: %s" % (type(e).__name__, e)}
        if isinstance(extras.get(name), dict):
            extras[name]["wall_s"] = round(time.perf_counter() - t0, 1)
    if rank == 0:
        if rec is not None:
            line = {"metric": METRIC, "value": rec["value"], "unit": UNIT, "n_gpus": world, "steps": steps,
                    "warmup": warmup, "ms_per_step": rec["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                    "vs_baseline": None, "dtype": "u8 text / int32 ids / f32 scores", "data": "synthetic"}
            for k in ("config", "workload_detail", "input_gb_per_s", "ids_per_s", "general_tier_sentences_per_step",
                      "viterbi_tier_sentences_per_step", "stage_ms_per_step", "roofline", "clocks", "gpu_launches",
                      "e2e", "parity", "cpu_baseline"):
                if k in rec:
                    line[k] = rec[k]
            if extras:
                line["configs"] = extras
            if failed:
                line["parity_failure"] = failed
            print(json.dumps(line), file=RESULT, flush=True)
        else:
            print(json.dumps({"metric": METRIC, "value": None, "parity_failure": failed}), file=RESULT, flush=True)
    if world > 1:
        ctx.dist.destroy_process_group()
    if failed:
        print("bench.py: PARITY FAILURE: " + failed, file=sys.stderr)
        sys.exit(3)


RESULT = sys.stdout


def main():
    global RESULT
    args = parse_args()
    # stdout carries exactly one JSON line: whatever else writes to fd 1 (NCCL prints its version
    # banner there with printf) goes to stderr instead
    sys.stdout.flush()
    RESULT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
