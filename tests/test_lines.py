"""Line splitting + spm_encode-compatible id output."""
import os
import sys

import numpy as np
import pytest

from conftest import ROOT, checker, model_bytes

sys.path.insert(0, os.path.join(ROOT, "tools"))


def getline_split(buf):
    """What a std::getline loop yields (src/filesystem.cc:42-44)."""
    lines = buf.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    return lines


def test_format_ids_matches_join():
    import spm_encode as t
    rng = np.random.default_rng(3)
    for _ in range(20):
        lens = rng.integers(0, 5, size=rng.integers(0, 30))
        off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        ids = rng.integers(0, 32000, size=int(off[-1])).astype(np.int32)
        want = b"".join(b" ".join(b"%d" % x for x in ids[off[k]:off[k + 1]]) + b"\n" for k in range(len(off) - 1))
        assert t.format_ids(ids, off) == want


@pytest.mark.gpu
@pytest.mark.parametrize("buf", [b"", b"\n", b"abc", b"abc\n", b"abc\ndef", b"a\r\nb\r\n", b"\n\nx\n\n", b"x" * 5000 + b"\n" + b"y" * 4095,
                                 b"\n" * 9000])
def test_split_lines_small(spm_b200, buf):
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(model_bytes("botchan_uni8k"))
    text, off = sp.split_lines(buf)
    want = getline_split(buf)
    assert len(off) - 1 == len(want)
    assert [bytes(text[off[i]:off[i + 1]]) for i in range(len(want))] == want


@pytest.mark.gpu
def test_encode_file_buffer_matches_reference(spm_b200):
    """botchan.txt as a file (CRLF line ends, config 1): ids per line = the reference's Encode per getline line."""
    import gzip
    import spm_encode as t
    mb = model_bytes("botchan_uni8k")
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    with gzip.open(os.path.join(ROOT, "tests", "golden", "data", "corpus_en_botchan.txt.gz"), "rb") as f:
        raw = f.read()
    lines = getline_split(raw)
    text, off = sp.split_lines(raw)
    assert len(off) - 1 == len(lines) == 4288
    assert bytes(text) == b"".join(lines)
    ids, io = sp.encode_file_buffer(raw)
    chk, _ = checker(mb)
    wi, wo = chk.encode_batch(np.frombuffer(b"".join(lines), dtype=np.uint8), off)
    assert np.array_equal(io, wo) and np.array_equal(ids, wi)
    out = t.format_ids(ids, io)
    assert out.count(b"\n") == 4288
    assert out.split(b"\n")[0] == b" ".join(b"%d" % x for x in wi[wo[0]:wo[1]])
