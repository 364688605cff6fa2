"""A/B builds of libspm_cuda.so: same sources, other build-time knobs (-D...), written next to the
product library as lib/variants/libspm_cuda_<name>.so (git-ignored; they travel to the GPU box).
usage: build_variants.py name1:-DX=1,-DY=2 name2:-DZ=3 ...   (bench.py picks one up through
SPM_B200_LIB_VARIANT=<path>)."""
import importlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import spm_b200_pkg  # noqa: E402

b = spm_b200_pkg.load()._build
out_dir = os.path.join(os.path.dirname(b.LIB), "variants")
os.makedirs(out_dir, exist_ok=True)
procs = []
for spec in sys.argv[1:]:
    name, _, flags = spec.partition(":")
    out = os.path.join(out_dir, "libspm_cuda_%s.so" % name)
    cmd = [os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")] + b.NVCC_FLAGS + [f for f in flags.split(",") if f] + \
        [os.path.join(b.CSRC, s) for s in b.SOURCES] + ["-o", out]
    procs.append((name, subprocess.Popen(cmd)))
for name, p in procs:
    if p.wait():
        raise SystemExit("variant %s failed to build" % name)
    print("built", name)
