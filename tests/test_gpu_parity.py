"""Parity of the CUDA path (through the C ABI) with the oracle, on a real GPU.

Bit-exact ids and CSR offsets are required (integer/index work).  The checker
is the unmodified reference compiled into oracle/_ref when that library is
present (it travels with the snapshot), else the CPU restatement.
"""
import numpy as np
import pytest

from conftest import EDGE_SENTENCES, checker, model_bytes

pytestmark = pytest.mark.gpu

CASES = [
    ("botchan_uni8k", "botchan"),      # config 1
    ("botchan_bpe4k", "botchan"),
    ("botchan_uni2k_ud", "botchan"),   # user-defined + control symbols
    ("botchan_bpe2k_ud", "botchan"),
    ("syn_uni32k", "syn_en"),          # config 2 (sample)
    ("syn_uni32k", "syn_dirty"),
    ("syn_bpe32k", "syn_en"),          # config 3 (sample)
    ("syn_bpe32k", "syn_dirty"),
    ("wagahai_uni", "wagahai"),        # config 4
    ("ref_ja_8k", "wagahai"),
    ("ref_en_1k", "botchan"),
    ("syn_uni32k", "wagahai"),         # mostly-unknown text, long unk runs
    ("syn_bpe32k", "wagahai"),
    ("wagahai_uni", "syn_dirty"),
]


def ref_encode(chk, text, off):
    try:
        return chk.encode_batch(text, off, threads=8)
    except TypeError:
        return chk.encode_batch(text, off)


def assert_same(got, want, text, off):
    gi, go = got
    wi, wo = want
    if not np.array_equal(go, wo):
        bad = np.nonzero(np.diff(go) != np.diff(wo))[0]
        i = int(bad[0])
        raise AssertionError("sentence %d: %d ids, want %d; text=%r" % (
            i, go[i + 1] - go[i], wo[i + 1] - wo[i], bytes(text[off[i]:off[i + 1]])[:200]))
    if not np.array_equal(gi, wi):
        k = int(np.nonzero(gi != wi)[0][0])
        i = int(np.searchsorted(wo, k, side="right") - 1)
        raise AssertionError("sentence %d differs: got %s want %s; text=%r" % (
            i, gi[wo[i]:wo[i + 1]][:40], wi[wo[i]:wo[i + 1]][:40], bytes(text[off[i]:off[i + 1]])[:200]))


@pytest.mark.parametrize("model,corpus", CASES)
def test_corpus_parity(spm_b200, corpora, model, corpus):
    mb = model_bytes(model)
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    chk, _ = checker(mb)
    text, off = corpora[corpus]
    assert_same(sp.encode_packed(text, off), ref_encode(chk, text, off), text, off)


@pytest.mark.parametrize("model", ["botchan_uni8k", "botchan_bpe4k", "botchan_uni2k_ud", "botchan_bpe2k_ud",
                                   "syn_uni32k", "syn_bpe32k", "wagahai_uni", "ref_ja_8k"])
def test_edge_sentences(spm_b200, corpora, model):
    mb = model_bytes(model)
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    chk, _ = checker(mb)
    text, off = corpora["edge"]
    assert_same(sp.encode_packed(text, off), ref_encode(chk, text, off), text, off)
    # single-sentence API
    for s in EDGE_SENTENCES[:40]:
        assert sp.EncodeAsIds(s) == chk.encode(s)


@pytest.mark.parametrize("model", ["botchan_uni8k", "botchan_bpe4k"])
@pytest.mark.parametrize("opts", ["bos", "eos", "reverse", "bos:eos", "reverse:bos:eos", "bos:eos:reverse",
                                  "eos:reverse:bos:reverse:eos"])
def test_extra_options(spm_b200, corpora, model, opts):
    mb = model_bytes(model)
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    sp.SetEncodeExtraOptions(opts)
    chk, _ = checker(mb)
    chk.set_encode_extra_options(opts)
    text, off = corpora["edge"]
    assert_same(sp.encode_packed(text, off), ref_encode(chk, text, off), text, off)
    text, off = corpora["botchan"]
    assert_same(sp.encode_packed(text, off), ref_encode(chk, text, off), text, off)


@pytest.mark.parametrize("model", ["botchan_uni8k", "botchan_bpe4k", "syn_bpe32k"])
def test_set_vocabulary(spm_b200, corpora, model):
    """SetVocabulary -> UNUSED pieces (unigram: skipped; BPE: resegmented through rev_merge)."""
    from oracle import ref
    if not ref.available():
        pytest.skip("needs oracle/_ref (SetVocabulary lives in the reference processor)")
    mb = model_bytes(model)
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    chk = ref.RefProcessor(mb)
    n = sp.GetPieceSize()
    vocab = [sp.IdToPiece(i) for i in range(n) if i % 3 != 0]
    sp.SetVocabulary(vocab)
    chk.set_vocabulary(vocab)
    text, off = corpora["syn_en" if model.startswith("syn") else "botchan"]
    text, off = text[: off[3000]], off[:3001]
    assert_same(sp.encode_packed(text, off), ref_encode(chk, text, off), text, off)
    sp.ResetVocabulary()
    chk.set_vocabulary(None)
    assert_same(sp.encode_packed(text, off), ref_encode(chk, text, off), text, off)


def test_device_memory_space(spm_b200, corpora):
    import torch
    mb = model_bytes("syn_uni32k")
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    text, off = corpora["syn_en"]
    want = sp.encode_packed(text, off)
    dt = torch.from_numpy(text).cuda()
    do = torch.from_numpy(off).cuda()
    ids = torch.empty(int(want[1][-1]) + 10, dtype=torch.int32, device="cuda")
    ido = torch.empty(len(off), dtype=torch.int64, device="cuda")
    sp.encode_device(dt, do, ids, ido)
    torch.cuda.synchronize()
    assert np.array_equal(ido.cpu().numpy(), want[1])
    assert np.array_equal(ids[: int(want[1][-1])].cpu().numpy(), want[0])
    # capacity too small -> RESOURCE_EXHAUSTED with the needed size in id_offsets[n]
    small = torch.empty(100, dtype=torch.int32, device="cuda")
    with pytest.raises(spm_b200.StatusError) as e:
        sp.encode_device(dt, do, small, ido)
    assert e.value.code == 8
    assert int(ido[-1].item()) == int(want[1][-1])


def test_empty_batch_and_errors(spm_b200):
    sp = spm_b200.SentencePieceProcessor()
    with pytest.raises(spm_b200.StatusError):
        sp.EncodeAsIds("x")  # model not loaded
    with pytest.raises(spm_b200.StatusError):
        sp.LoadFromSerializedProto(b"garbage")
    sp.LoadFromSerializedProto(model_bytes("botchan_uni8k"))
    ids, off = sp.encode_batch([])
    assert len(ids) == 0 and off.tolist() == [0]
    assert sp.EncodeAsIds("") == []
    with pytest.raises(spm_b200.StatusError):
        sp.SetEncodeExtraOptions("foo")


@pytest.mark.parametrize("model,corpus", [("botchan_uni8k", "botchan"), ("botchan_bpe4k", "botchan"),
                                          ("ref_ja_8k", "wagahai"), ("syn_bpe32k", "syn_dirty")])
def test_long_lines(spm_b200, corpora, model, corpus):
    """Lines around the tier boundaries (4096 and 65536 raw bytes) among short ones."""
    mb = model_bytes(model)
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    chk, _ = checker(mb)
    text, off = corpora[corpus]
    reps = -(-(2 << 20) // len(text))
    off = np.concatenate([[0]] + [off[1:] + r * len(text) for r in range(reps)]).astype(np.int64)
    text = np.tile(text, reps)
    rng = np.random.default_rng(5)
    n = len(off) - 1
    # glue runs of consecutive sentences into long lines
    cuts = [0]
    targets = [4095, 4096, 4097, 5000, 9000, 20000, 33000, 65535, 65536, 70000, 150000]
    k = 0
    i = 0
    while i < n:
        if k < 200 and rng.random() < 0.05:
            want = targets[k % len(targets)] if k < 2 * len(targets) else int(rng.integers(4096, 40000))
            k += 1
            j = i
            while j < n and off[j] - off[i] < want:
                j += 1
            # exact byte length where the text allows it (cut inside a sentence: keep UTF-8 whole)
            end = min(int(off[i]) + want, int(off[n]))
            while end < int(off[n]) and (text[end] & 0xC0) == 0x80:
                end += 1
            cuts.append(end)
            i = int(np.searchsorted(off, end, side="left"))
            if i <= n and i < len(off) and off[i] != end:
                cuts.append(int(off[i]))
        else:
            i += 1
            cuts.append(int(off[i]))
    cuts = np.unique(np.asarray(cuts, dtype=np.int64))
    lens = np.diff(cuts)
    assert (lens >= 4096).sum() >= 10 and (lens >= 65536).sum() >= 1
    got = sp.encode_packed(text, cuts)
    assert_same(got, ref_encode(chk, text, cuts), text, cuts)
    st = sp.last_stats()
    # only the >= 64 KiB lines (and the rare flagged ones) may take the general tier
    # (dirty text: an over-long word sends the whole line to the general tier; BPE: a short leftover
    # list goes to the general tier as a whole -- option bpe_route)
    if corpus != "syn_dirty" and "bpe" not in model:
        assert st["fallback_sentences"] < (lens >= 4096).sum()
    if "bpe" in model:          # the merge kernel itself on the same lines
        sp.set_option("bpe_route", 0)
        assert_same(sp.encode_packed(text, cuts), ref_encode(chk, text, cuts), text, cuts)
        assert sp.last_stats()["fallback_sentences"] < (lens >= 4096).sum() or corpus == "syn_dirty"


@pytest.mark.parametrize("model", ["syn_uni32k", "syn_bpe32k", "botchan_uni8k"])
def test_word_scan_shapes(spm_b200, corpora, model):
    """k_memo_scan_words on the shapes its CPU replay covers, end to end against the reference:
    words of 1 / 14 / 15 / 24 / 25 bytes, empty and one-byte sentences, leading / trailing /
    double spaces, control and non-ASCII bytes, warps with more words than the shared word list
    holds (per-lane fallback), warps whose text does not fit the staging buffer, sentences
    at every alignment of the 8-byte sweep."""
    mb = model_bytes(model)
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    chk, _ = checker(mb)
    rng = np.random.default_rng(11)
    text, off = corpora["syn_en"]
    words = bytes(text[: off[400]]).split()
    # (runs of 26 .. 1000 word bytes: the scan keeps 8-bit word ends and must flag them by its chunk rule)
    words += [b"v" * n for n in (26, 31, 32, 33, 39, 40, 41, 47, 48, 49, 255, 256, 257, 258, 264, 270, 280, 300, 511,
                                 512, 513, 520, 768, 1000)]
    words += [b"a", b"I", b"x" * 14, b"y" * 15, b"internationalisations", b"z" * 24, b"w" * 25, b"q" * 60,
              b"counter-revolutionary.", b"Mr.", b"(sic)", b"~", b"caf\xc3\xa9", b"na\xc3\xafve", b"\xef\xbd\x81"]
    lines = []
    for i in range(6000):
        r = rng.random()
        if r < 0.05:
            lines.append(b"")
        elif r < 0.08:
            lines.append(bytes([int(rng.integers(0x20, 0x7F))]))
        elif r < 0.11:     # many one-letter words: more than 640 words in a warp
            lines.append(b" ".join([b"a", b"b", b"I", b"o"][int(k)] for k in rng.integers(0, 4, size=int(rng.integers(30, 70)))))
        elif r < 0.12:     # a long line: the warp's text does not fit its staging buffer
            lines.append(b" ".join(words[int(k)] for k in rng.integers(0, 400, size=int(rng.integers(700, 1200)))))
        else:
            k = int(rng.integers(1, 24))
            ws = [words[int(j)] for j in rng.integers(0, len(words), size=k)]
            s = b" ".join(ws)
            q = rng.random()
            if q < 0.03:
                s = b" " + s
            elif q < 0.06:
                s = s + b" "
            elif q < 0.09:
                s = s.replace(b" ", b"  ", 1)
            elif q < 0.11:
                s = s.replace(b" ", b"\t", 1)
            elif q < 0.12:
                s = s + b"\r"
            lines.append(s)
    from tools import synth
    t2, o2 = synth.pack_lines(lines)
    want = ref_encode(chk, t2, o2)
    assert_same(sp.encode_packed(t2, o2), want, t2, o2)
    import torch
    dt, do = torch.from_numpy(t2).cuda(), torch.from_numpy(o2).cuda()
    ids = torch.empty(int(want[1][-1]) + 8, dtype=torch.int32, device="cuda")
    ido = torch.empty(len(o2), dtype=torch.int64, device="cuda")
    sp.encode_device(dt, do, ids, ido)
    torch.cuda.synchronize()
    assert_same((ids[: int(want[1][-1])].cpu().numpy(), ido.cpu().numpy()), want, t2, o2)
