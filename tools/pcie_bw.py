"""Pinned host <-> device copy bandwidth of this box (one direction, then both at once)."""
import torch

n = 1 << 30
h1 = torch.empty(n, dtype=torch.uint8).pin_memory()
h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
d1 = torch.empty(n, dtype=torch.uint8, device="cuda")
d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def h2d():
    d1.copy_(h1, non_blocking=True)


def d2h():
    h2.copy_(d2, non_blocking=True)


def both():
    with torch.cuda.stream(s1):
        d1.copy_(h1, non_blocking=True)
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)


for name, fn in (("h2d", h2d), ("d2h", d2h), ("both", both)):
    ms = timed(fn)
    print("%s: %.2f ms per GiB -> %.1f GB/s%s" % (name, ms, n / ms / 1e6, " per direction" if name == "both" else ""))
