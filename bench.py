#!/usr/bin/env python3
"""bench.py -- batch EncodeAsIds throughput on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One "step" = one pass of spm_cuda_encode_batch over one batch of synthetic
sentences (config 2 of BASELINE.json: unigram-32k, nmt_nfkc, 10 M English-like
sentences of ~100 B per GPU; sentence-sharded, no collective, weak scaling).
`value` times the device-resident call (CUDA events on the launching stream);
`e2e` times the same call with HOST (pinned) buffers, H2D/D2H inside the timed
region.  `--impl reference` times the unmodified reference compiled into
oracle/_ref (all host threads) on a bounded sample of the same workload.
"""
import argparse
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


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="syn_uni32k")
    ap.add_argument("--sentences", type=int, default=10_000_000, help="sentences per GPU per step")
    ap.add_argument("--corpus", default="syn_en", choices=["syn_en", "syn_dirty", "wagahai"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="sentences of the cpu_baseline sample (0: auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def model_bytes(name):
    path = name if os.path.exists(name) else os.path.join(MODELS, name + ".model")
    with open(path, "rb") as f:
        return f.read()


def make_corpus(kind, n, first):
    from tools import synth
    if kind == "syn_en":
        return synth.EnglishCorpus().generate(n, first=first)
    if kind == "syn_dirty":
        return synth.EnglishCorpus(dirty=True).generate(n, first=first)
    return synth.ResampledCorpus().generate(n, first=first)


def workload_name(args):
    return {"syn_en": "config2: %s, %d synthetic English-like sentences (~100 B) per GPU",
            "syn_dirty": "config2-dirty: %s, %d synthetic sentences with NFKC/whitespace/malformed noise per GPU",
            "wagahai": "config4: %s, %d wagahai-style CJK lines per GPU"}[args.corpus] % (args.model, args.sentences)


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


def ref_processor(mb):
    """The CPU arm: the unmodified reference (oracle/_ref) when its library is
    present, else the CPU restatement (oracle port; single-threaded)."""
    from oracle import port, ref
    if ref.available():
        return ref.RefProcessor(mb), "reference"
    return port.OracleProcessor(mb), "port"


def time_cpu(chk, kind, text, off, threads):
    t0 = time.perf_counter()
    if kind == "reference":
        n_ids = chk.encode_count(text, off, threads=threads)
    else:
        n_ids = len(chk.encode_batch(text, off)[0])
    return time.perf_counter() - t0, n_ids


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
        time_cpu(chk, kind, text, off, threads)
    t = 0.0
    for _ in range(args.steps):
        dt, n_ids = time_cpu(chk, kind, text, off, threads)
        t += dt
    v = n * args.steps / t
    sample = "first %d sentences of the workload per step (%d B, %d ids), in-process Encode(ids) loop, no file I/O" % (
        n, int(off[-1]), n_ids)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int32 ids, f32 scores",
            "data": "synthetic", "config": {"workload": workload_name(args), "sample": sample},
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


STAGE_KERNEL = {"encode": "k_encode_fast / k_encode_bpe_fast (general tier kernels for models without a fast tier)",
                "normalize": "k_normalize_fast", "memo_collect": "k_memo_scan_words (+ k_memo_scan for the listed sentences)", "memo_words": "k_memo_words_*",
                "memo_emit": "k_memo_emit", "compact": "k_scan_* + k_compact"}


def run_b200(args):
    import torch
    import torch.distributed as dist
    import spm_b200_pkg

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this framework has no CPU path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"      # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=dev)
    numa = bind_to_gpu_numa_node(local) if world > 1 else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    pkg = spm_b200_pkg.load()
    mb = model_bytes(args.model)
    sp = pkg.SentencePieceProcessor(device=local)
    sp.LoadFromSerializedProto(mb)

    n = args.sentences
    text, off = make_corpus(args.corpus, n, rank * n)   # this rank's shard (contiguous sentence range)
    in_bytes = int(off[-1])
    d_text = torch.from_numpy(text).to(dev)
    d_off = torch.from_numpy(off).to(dev)
    d_ids = torch.empty(in_bytes + 64, dtype=torch.int32, device=dev)   # <= one id per input byte
    d_idoff = torch.empty(n + 1, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream(dev)

    def step_device():
        sp.encode_device(d_text, d_off, d_ids, d_idoff, stream=stream.cuda_stream)
        return sp.last_stats()

    st = None
    for _ in range(max(args.warmup, 3)):
        st = step_device()
    n_ids = st["ids"]
    alg_bytes = in_bytes + 8 * n + 4 * n_ids + 8 * n   # per sentence in + 8 + 4 ids + 8

    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kern_ms, kern_launches, launches = 0.0, 0, 0
    stage_ms = {}
    barrier()
    with ClockSampler(local) as clk:
        ev0.record(stream)
        for _ in range(args.steps):
            st = step_device()
            launches += st["launches"]
            for k, v in st["stage_ms"].items():
                stage_ms[k] = stage_ms.get(k, 0.0) + v
        ev1.record(stream)
        barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = clk.summary()
    # the dominant stage of the step: one kernel per call (two when the batch holds lines >= 4 KiB)
    dom = max(stage_ms, key=stage_ms.get)
    kern_ms, kern_launches = stage_ms[dom], args.steps

    # ---- e2e: same call, HOST (pinned) buffers, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        h_text = torch.from_numpy(text).pin_memory()
        h_off = torch.from_numpy(off).pin_memory()
        h_ids = torch.empty(n_ids + 64, dtype=torch.int32).pin_memory()
        h_idoff = torch.empty(n + 1, dtype=torch.int64).pin_memory()
        for _ in range(2):
            sp.encode_pinned(h_text, h_off, h_ids, h_idoff, stream=stream.cuda_stream)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(args.steps):
            sp.encode_pinned(h_text, h_off, h_ids, h_idoff, stream=stream.cuda_stream)
        e1.record(stream)
        barrier()
        e2e_ms = e0.elapsed_time(e1)
        assert int(h_idoff[n]) == n_ids
        e2e = (e2e_ms, in_bytes + 8 * (n + 1), 4 * n_ids + 8 * (n + 1))

    # ---- max over ranks
    t = torch.tensor([ms, e2e[0] if e2e else 0.0], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(n), float(in_bytes), float(n_ids)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_all, e2e_all = t.tolist()
    n_all, bytes_all, ids_all = tot.tolist()

    if rank == 0:
        peak, peak_src = measured_peak()
        # dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, one launch, from the
        # ncu --set full capture of this very workload (k_memo_scan_words: 1.44 + 0.80 GB,
        # profiles/r1_v23_summary.md / r1_v21_ncu_full_keys.txt); null otherwise
        traffic = None
        if (dom == "memo_collect" and args.model == "syn_uni32k" and args.corpus == "syn_en"
                and n == 10_000_000):
            traffic = 2.234e9
        k_avg_ms = kern_ms / max(kern_launches, 1)
        alg_per_launch = alg_bytes * args.steps / max(kern_launches, 1)
        achieved = alg_per_launch / (k_avg_ms * 1e-3) / 1e9 if k_avg_ms > 0 else 0.0
        line = {"metric": METRIC, "value": n_all * args.steps / (ms_all * 1e-3), "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_all / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8 text / int32 ids / f32 scores", "data": "synthetic",
                "config": {"workload": workload_name(args), "sentences_per_gpu": n, "input_bytes_per_gpu": in_bytes,
                           "ids_per_gpu": n_ids, "l2": "inputs (%.2f GB) larger than L2; no flush" % (in_bytes / 1e9),
                           "parallelism": "sentence-sharded x%d, no collective" % world,
                           "numa_node_rank0": numa},
                "input_gb_per_s": bytes_all * args.steps / (ms_all * 1e-3) / 1e9,
                "ids_per_s": ids_all * args.steps / (ms_all * 1e-3),
                "general_tier_sentences_per_step": st["fallback_sentences"],
                "viterbi_tier_sentences_per_step": st["viterbi_tier_sentences"],
                "stage_ms_per_step": {k: round(v / args.steps, 4) for k, v in stage_ms.items()},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": traffic, "kernel": STAGE_KERNEL.get(dom, dom),
                             "kernel_ms_per_launch": k_avg_ms, "kernel_launches": kern_launches,
                             "kernel_share_of_step": kern_ms / ms if ms > 0 else None,
                             "algorithmic_bytes_per_launch": alg_per_launch, "peak_source": peak_src},
                "clocks": clocks, "gpu_launches": launches}
        if e2e:
            line["e2e"] = {"value": n_all * args.steps / (e2e_all * 1e-3), "unit": UNIT,
                           "h2d_bytes_per_step": e2e[1], "d2h_bytes_per_step": e2e[2],
                           "ms_per_step": e2e_all / args.steps}
        if world == 1 and not args.no_cpu_baseline:
            chk, kind = ref_processor(mb)
            threads = host_threads() if kind == "reference" else 1
            m = args.cpu_sample or min(n, 125_000 * threads)
            dt, cids = time_cpu(chk, kind, text[: off[m]], off[: m + 1], threads)
            line["cpu_baseline"] = {"value": m / dt, "unit": UNIT, "cores": threads, "kind": kind,
                                    "sample": "first %d sentences of the workload (%d B, %d ids), %.1f s wall, "
                                              "in-process Encode(ids) loop" % (m, int(off[m]), cids, dt)}
        print(json.dumps(line), file=RESULT, flush=True)
    if world > 1:
        dist.destroy_process_group()


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
