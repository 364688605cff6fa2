"""spm_cuda_split_lines_noalloc (include/spm_cuda.h): the allocation-free variant of line splitting,
called through the C ABI, against the existing entry point (spm_cuda_split_lines, DEVICE form) on
two calls in a row -- integer / byte outputs, so the bar is bit-exact.  The second input is
shorter than the first: stale entries of the shared static scratch would show.  Collected last on
purpose (file name)."""
import ctypes

import numpy as np
import pytest

from conftest import model_bytes

pytestmark = pytest.mark.gpu

LIMIT = 256 << 20


def buffers():
    from tools import synth
    bot = b"\n".join(synth.read_gz_lines("corpus_en_botchan.txt.gz")) + b"\n"       # CRLF file, 4,288 lines
    wag = b"\n".join(synth.read_gz_lines("corpus_ja_wagahai.txt.gz"))              # no newline at the end
    rng = np.random.default_rng(7)
    noise = rng.integers(0, 256, size=300_000, dtype=np.uint8)
    noise[rng.integers(0, len(noise), size=20_000)] = 10
    return [bot * 3 + wag, wag[: len(wag) // 3] + b"\n\n\nlast line without newline", noise.tobytes(),
            b"\n", b"x", b"\n\n\n", b"a\nb"]


def run(sp, lib, h, buf, noalloc):
    import torch
    d_buf = torch.from_numpy(np.frombuffer(buf, dtype=np.uint8).copy()).cuda()
    cap = len(buf) + 1
    d_text = torch.zeros(len(buf) + 16, dtype=torch.uint8, device="cuda")
    d_off = torch.full((cap + 1,), -7, dtype=torch.int64, device="cuda")
    n = sp.split_lines_device(d_buf, d_text, d_off, noalloc=noalloc)
    torch.cuda.synchronize()
    off = d_off[: n + 1].cpu().numpy()
    return n, off, d_text[: int(off[-1])].cpu().numpy()


def test_noalloc_matches_existing_entry_point(spm_b200):
    import torch
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(model_bytes("botchan_uni8k"))
    lib = spm_b200.lib()
    bufs = buffers()
    want = [run(sp, lib, sp._h, b, False) for b in bufs]
    for rep in range(2):                       # two passes: every call follows a call with another input
        for b, (wn, woff, wtext) in zip(bufs, want):
            n, off, text = run(sp, lib, sp._h, b, True)
            assert n == wn
            assert np.array_equal(off, woff)
            assert np.array_equal(text, wtext)
    # against the definition as well (std::getline over the buffer)
    n, off, text = run(sp, lib, sp._h, bufs[1], True)
    lines = bufs[1].split(b"\n")
    if lines[-1] == b"":
        lines.pop()
    assert n == len(lines) and text.tobytes() == b"".join(lines)
    assert np.array_equal(np.diff(off), [len(x) for x in lines])


def test_noalloc_allocates_nothing_and_checks_its_limit(spm_b200):
    import torch
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(model_bytes("botchan_uni8k"))
    L = spm_b200.lib()
    buf = buffers()[0]
    d_buf = torch.from_numpy(np.frombuffer(buf, dtype=np.uint8).copy()).cuda()
    d_text = torch.empty(len(buf) + 16, dtype=torch.uint8, device="cuda")
    d_off = torch.empty(len(buf) + 2, dtype=torch.int64, device="cuda")
    n0 = sp.split_lines_device(d_buf, d_text, d_off, noalloc=True)       # first use: module load, symbol lookup
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        assert sp.split_lines_device(d_buf, d_text, d_off, noalloc=True) == n0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, "device memory changed across allocation-free calls"
    # NULL model handle: the current device
    n = ctypes.c_int64(0)
    rc = L.spm_cuda_split_lines_noalloc(None, d_buf.data_ptr(), d_buf.numel(), d_text.data_ptr(), d_off.data_ptr(),
                                        d_off.numel() - 1, ctypes.byref(n), None)
    assert rc == 0 and n.value == n0
    # capacity too small: RESOURCE_EXHAUSTED with the line count
    rc = L.spm_cuda_split_lines_noalloc(sp._h, d_buf.data_ptr(), d_buf.numel(), d_text.data_ptr(), d_off.data_ptr(),
                                        10, ctypes.byref(n), None)
    assert rc == 8 and n.value == n0
    # beyond the static scratch: refused before anything is touched
    rc = L.spm_cuda_split_lines_noalloc(sp._h, d_buf.data_ptr(), LIMIT + 1, d_text.data_ptr(), d_off.data_ptr(),
                                        d_off.numel() - 1, ctypes.byref(n), None)
    assert rc == 11
    msg = L.spm_cuda_last_error().decode()
    assert str(LIMIT) in msg and "exceeds" in msg
    # empty buffer
    rc = L.spm_cuda_split_lines_noalloc(sp._h, None, 0, None, d_off.data_ptr(), 4, ctypes.byref(n), None)
    assert rc == 0 and n.value == 0 and int(d_off[0].item()) == 0


def test_noalloc_at_the_limit(spm_b200):
    """Exactly 256 MiB: all 65,536 tiles of the static scratch."""
    import torch
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(model_bytes("botchan_uni8k"))
    big = torch.full((LIMIT,), 97, dtype=torch.uint8, device="cuda")
    big[999::1000] = 10
    big[-1] = 98                                     # unterminated last line
    cap = LIMIT // 1000 + 2
    outs = []
    for noalloc in (False, True):
        d_text = torch.zeros(LIMIT + 16, dtype=torch.uint8, device="cuda")
        d_off = torch.full((cap + 1,), -1, dtype=torch.int64, device="cuda")
        n = sp.split_lines_device(big, d_text, d_off, noalloc=noalloc)
        torch.cuda.synchronize()
        outs.append((n, d_off[: n + 1].clone(), d_text[: int(d_off[n].item())].clone()))
    assert outs[0][0] == outs[1][0] == LIMIT // 1000 + 1
    assert torch.equal(outs[0][1], outs[1][1]) and torch.equal(outs[0][2], outs[1][2])
    assert int(outs[1][1][-1].item()) == LIMIT - LIMIT // 1000
