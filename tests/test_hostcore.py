"""The kernels' __host__ __device__ arithmetic (spm_core.cuh exact tier,
spm_fast.cuh fast tier, host model loader) compiled for the CPU by
tests/host_core.cc and compared with the reference fixtures / the oracle.
This is how every kernel change is parity-checked in the GPU-less container
before it runs on a B200; the shipped library has no host encode path."""
import ctypes
import os

import numpy as np
import pytest

from conftest import EDGE_SENTENCES, GOLDEN, checker, model_bytes
from oracle import port

import sys
sys.path.insert(0, GOLDEN)
import make_fixtures as mf  # noqa: E402


@pytest.fixture(scope="module")
def ref_ids():
    return np.load(os.path.join(GOLDEN, "ref_ids.npz"))


def is_unigram(model):
    return "bpe" not in model


@pytest.mark.parametrize("model,corpus,first", mf.CASES)
def test_core_matches_reference_fixture(make_hostcore, corpora, ref_ids, model, corpus, first):
    text, off = corpora[corpus]
    first = min(first or len(off) - 1, 2000)
    text, off = text[: off[first]], off[: first + 1]
    wo = ref_ids["%s__%s__off" % (model, corpus)][: first + 1]
    wi = ref_ids["%s__%s__ids" % (model, corpus)][: wo[-1]]
    hc = make_hostcore(model_bytes(model))
    # exact tier (2 capacities); 0 = fast tier (unigram / word-wise BPE where the model allows);
    # -1 = word memo in front of the fast tier
    tiers = [4, 16, 0, -1]
    for capc in tiers:
        gi, go, st = hc.encode_batch(text, off, capc=capc)
        assert np.array_equal(go, wo) and np.array_equal(gi, wi), (model, corpus, capc)
        if capc == 16:
            assert not (st & 4).any(), "exact tier overflowed its candidate list"


def test_core_near_tie_retry(make_hostcore):
    """More than 8 near-tied candidates at one position: the exact tier's first pass overflows
    (capacity 8 as in the library), the retry at min(ring, 255) reproduces the reference."""
    from conftest import NEAR_TIE_SENTENCES, near_tie_model
    from tools import synth
    mb = near_tie_model()
    text, off = synth.pack_lines(NEAR_TIE_SENTENCES)
    chk, _ = checker(mb)
    wi, wo = chk.encode_batch(text, off)
    hc = make_hostcore(mb)
    gi, go, st = hc.encode_batch(text, off, capc=8)
    assert np.array_equal(go, wo) and np.array_equal(gi, wi)
    assert (st & (1 << 23)).sum() >= 4, "the crafted sentences did not overflow the 8-candidate lists"
    assert not (st & 4).any()
    for capc in (0, -1):      # fast tier / word memo in front: flagged, served by the exact tier
        gi, go, _ = hc.encode_batch(text, off, capc=capc)
        assert np.array_equal(go, wo) and np.array_equal(gi, wi), capc


@pytest.mark.parametrize("model", ["ref_ja_8k", "wagahai_uni"])
def test_fast_tier_near_ties_in_line(make_hostcore, model, monkeypatch):
    """fast_unigram keeps near-tied candidates itself (side area) instead of flagging the sentence: on long
    Japanese lines -- where the tolerance, which grows with the running score, meets a near tie almost surely --
    nearly every line is served by the fast tier itself (status bit 24), with the reference's ids; with the side
    area off (round 1's behaviour) most of these lines are flagged (status bit 2) and re-run by the exact tier."""
    from tools import synth
    lines = [l for l in synth.read_gz_lines("corpus_ja_wagahai.txt.gz") if len(l) >= 1500]
    # longer ones still: pairs of lines glued together (up to ~40 KB)
    lines += [a + b for a, b in zip(lines[::2], lines[1::2])][:40]
    text, off = synth.pack_lines(lines)
    mb = model_bytes(model)
    chk, _ = checker(mb)
    wi, wo = chk.encode_batch(text, off)
    hc = make_hostcore(mb)
    gi, go, st = hc.encode_batch(text, off, capc=0)
    assert np.array_equal(go, wo) and np.array_equal(gi, wi)
    served = int(((st >> 24) & 1).sum())
    monkeypatch.setenv("HC_FAST_NO_SIDE", "1")
    gi, go, st0 = hc.encode_batch(text, off, capc=0)
    assert np.array_equal(go, wo) and np.array_equal(gi, wi)
    flagged0 = int(((st0 >> 2) & 1).sum())
    print("%s: %d lines, served in line %d, flagged without the side area %d" % (model, len(lines), served, flagged0))
    n = np.diff(off)
    for lo, hi in ((1500, 4096), (4096, 8192), (8192, 16384), (16384, 1 << 20)):
        m = (n >= lo) & (n < hi)
        print("   %6d..%7d bytes: %3d lines, %3d flagged without the side area, %3d not served in line" % (
            lo, hi, m.sum(), ((st0 >> 2) & 1)[m].sum(), m.sum() - ((st >> 24) & 1)[m].sum()))
    assert flagged0 >= 10, "these lines were expected to meet near ties"
    assert served >= len(lines) - max(2, len(lines) // 20)


@pytest.mark.parametrize("model", mf.EDGE_MODELS)
def test_core_edge_sentences(make_hostcore, corpora, ref_ids, model):
    text, off = corpora["edge"]
    hc = make_hostcore(model_bytes(model))
    for capc in [4, 0, -1]:
        gi, go, _ = hc.encode_batch(text, off, capc=capc)
        assert np.array_equal(go, ref_ids[model + "__edge__off"]), (model, capc)
        assert np.array_equal(gi, ref_ids[model + "__edge__ids"]), (model, capc)


@pytest.mark.parametrize("model", ["botchan_uni8k", "botchan_uni2k_ud", "ref_ja_8k", "botchan_bpe2k_ud"])
def test_normalizers_match_reference(make_hostcore, corpora, model):
    """normalize_sentence (general tier) and fast_normalize (fast tier) vs the reference strings."""
    norm = np.load(os.path.join(GOLDEN, "ref_norm.npz"))
    hc = make_hostcore(model_bytes(model))
    text, off = norm[model + "__text"], norm[model + "__off"]
    routed = 0
    for k, s in enumerate(EDGE_SENTENCES):
        want = bytes(text[off[k]:off[k + 1]])
        assert hc.normalize(s) == want, (model, s)
        f = hc.normalize_fast(s)
        if f is None:
            routed += 1
        else:
            assert f[0] == want, (model, s)
            assert f[1] >= len(want.decode("utf-8", "replace")), "nchars must be an upper bound"
    assert routed <= 4   # only the heavy-expansion inputs may leave the fast tier
    # and against the oracle on a dirty synthetic sample
    op = port.OracleProcessor(model_bytes(model))
    t, o = corpora["syn_dirty"]
    for i in range(0, 3000):
        s = bytes(t[o[i]:o[i + 1]])
        want = op.normalize(s)[0]
        assert hc.normalize(s) == want
        f = hc.normalize_fast(s)
        assert f is None or f[0] == want, s


@pytest.mark.parametrize("model", ["botchan_uni8k", "botchan_uni2k_ud", "ref_ja_8k", "botchan_bpe2k_ud"])
def test_piecewise_normalizer_matches(make_hostcore, corpora, model):
    """k_normalize_long's scheme (pieces cut at norm_find_cut positions, normalized independently,
    concatenated) gives the bytes of the sequential normalizer -- with tiny pieces so that cuts
    fall next to rules, user symbols, whitespace runs and malformed bytes."""
    hc = make_hostcore(model_bytes(model))
    lines = [s for s in EDGE_SENTENCES if s]
    for name, step, sep in (("wagahai", 40, b""), ("syn_dirty", 60, b"  "), ("botchan", 50, b" ")):
        t, o = corpora[name]
        for i in range(0, min(len(o) - 1, 1200), step):
            lines.append(sep.join(bytes(t[o[k]:o[k + 1]]) for k in range(i, i + step)))
    lines.append(b" ".join(EDGE_SENTENCES))
    lines.append(b"".join(EDGE_SENTENCES))
    served = 0
    for s in lines:
        want = hc.normalize(s)
        for chunk in (3, 16, 512):
            got = hc.normalize_chunked(s, chunk)
            if got is not None:
                served += 1
                assert got == want, (model, chunk, s[:80])
    assert served >= 2 * len(lines)


@pytest.mark.parametrize("opts", ["bos", "eos:reverse", "reverse:bos:eos", "eos:reverse:bos:reverse:eos"])
def test_extra_options(make_hostcore, corpora, opts):
    mb = model_bytes("botchan_uni8k")
    chk, _ = checker(mb)
    chk.set_encode_extra_options(opts)
    text, off = corpora["edge"]
    wi, wo = chk.encode_batch(text, off)
    gi, go, _ = make_hostcore(mb).encode_batch(text, off, opts=port.pack_options(opts), capc=0)
    assert np.array_equal(go, wo) and np.array_equal(gi, wi)


def test_set_piece_types(make_hostcore, corpora):
    """SetVocabulary semantics (UNUSED pieces) through the fast view rebuild."""
    mb = model_bytes("botchan_uni8k")
    op = port.OracleProcessor(mb)
    t = op.piece_types()
    t[(np.arange(len(t)) % 3 == 0) & (t == 1)] = 5
    op.set_piece_types(t)
    hc = make_hostcore(mb)
    hc.set_piece_types(t)
    text, off = corpora["botchan"]
    text, off = text[: off[800]], off[:801]
    wi, wo = op.encode_batch(text, off)
    for capc in (0, 4, -1):
        gi, go, _ = hc.encode_batch(text, off, capc=capc)
        assert np.array_equal(go, wo) and np.array_equal(gi, wi), capc


def test_fast_tier_eligibility(make_hostcore):
    """Which fixture models run on the fast tiers (the rest take the general tier)."""
    want = {"botchan_uni8k": 1, "syn_uni32k": 1, "wagahai_uni": 1, "ref_ja_8k": 1, "ref_en_1k": 1, "botchan_uni2k_ud": 1,
            "botchan_bpe4k": 2, "syn_bpe32k": 2, "botchan_bpe2k_ud": 0}   # user-defined symbols freeze BPE symbols
    for name, flags in want.items():
        assert make_hostcore(model_bytes(name)).fast_flags() & 3 == flags, name


def test_word_memo_serves_and_matches(make_hostcore, corpora):
    """The word memo (spm_memo.cuh) serves most sentences of word-delimited text and agrees
    with the reference; CJK lines without spaces fall through to the Viterbi tier."""
    for model, corpus, lo in [("syn_uni32k", "syn_en", 0.95), ("syn_bpe32k", "syn_en", 0.95),
                              ("botchan_uni8k", "botchan", 0.9), ("botchan_bpe4k", "botchan", 0.9),
                              ("syn_uni32k", "syn_dirty", 0.0), ("ref_ja_8k", "wagahai", 0.0)]:
        mb = model_bytes(model)
        hc = make_hostcore(mb)
        assert hc.fast_flags() & 4, model
        text, off = corpora[corpus]
        m = min(3000, len(off) - 1)
        text, off = text[: off[m]], off[: m + 1]
        chk, _ = checker(mb)
        wi, wo = chk.encode_batch(text, off)
        gi, go, st = hc.encode_batch(text, off, capc=-1)
        assert np.array_equal(go, wo) and np.array_equal(gi, wi), (model, corpus)
        served = float(((st >> 20) & 1).mean())
        assert served >= lo, (model, corpus, served)
        plain = float(((st >> 23) & 1).mean())     # sentences keyed by memo_scan_plain
        if corpus == "syn_en":
            assert plain >= 0.9, (model, plain)


def test_word_scan_matches_sentence_scan(hostcore, corpora):
    """k_memo_scan_words' arithmetic (sw_chunk_masks, sw_word_slot; lanes replayed on the CPU)
    keys every word exactly as memo_scan_plain does: same word counts, same slots, no extra
    keys -- on clean text, on text with NFKC / control / malformed bytes, and on edge shapes."""
    rng = np.random.default_rng(7)
    edge = [b"", b"a", b" ", b"a b", b" a", b"a ", b"a  b", b"abcdefghijklmn", b"abcdefghijklmno", b"x" * 24, b"y" * 25,
            b"abcdefghijklmn abcdefghijklmnopqrstuvwx z", b"a\tb", b"caf\xc3\xa9 au lait", b"~ ! ~", b"\x7f", b"a \x1f",
            b"", b"", b"z", b"ab cd ef gh ij kl mn op qr st uv wx yz" * 3, b"q" * 60 + b" r"]
    for _ in range(300):    # random shapes around the 8-byte chunk borders
        k = int(rng.integers(0, 40))
        edge.append(bytes(rng.choice(np.frombuffer(b"ab  c.Z~\t\r\n\x0b", dtype=np.uint8), size=k).tolist()))
    etext = np.frombuffer(b"".join(edge), dtype=np.uint8)
    eoff = np.concatenate([[0], np.cumsum([len(e) for e in edge])]).astype(np.int64)
    cases = [(etext, eoff)]
    for name in ("syn_en", "syn_dirty", "botchan"):
        text, off = corpora[name]
        m = min(4000, len(off) - 1)
        cases.append((np.ascontiguousarray(text[: off[m]]), np.ascontiguousarray(off[: m + 1])))
    hostcore.hc_scan_words_check.restype = ctypes.c_int64
    hostcore.hc_scan_words_check.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_int,
                                             ctypes.c_void_p]
    plain = {0: [], 1: [], 2: []}
    for relax in (0, 1, 2):  # 1: a model that removes extra whitespace -- leading / trailing / repeated spaces only separate words; 2: tabs too
        for text, off in cases:
            for prefix in (1, 0):
                n_plain = ctypes.c_int64(0)
                rc = hostcore.hc_scan_words_check(text.ctypes.data, off.ctypes.data, len(off) - 1, prefix, relax,
                                                  ctypes.byref(n_plain))
                assert rc == 0, (rc, relax)
            plain[relax].append(n_plain.value)
    # clean text: all plain; "dirty": most (botchan lines end in \r: none); more of the dirty and of the edge shapes when relaxed
    assert plain[0][1] >= 3900 and 2000 < plain[0][2] < 3900, plain
    assert plain[1][1] == plain[0][1] and plain[1][2] > plain[0][2] + 50 and plain[1][0] > plain[0][0] + 5, plain
    assert plain[2][1] == plain[0][1] and plain[2][2] > plain[1][2] + 30 and plain[2][0] > plain[1][0] + 30, plain
    assert plain[0][3] == 0 and plain[2][3] >= 3900, plain      # botchan: CR at every line end


@pytest.mark.parametrize("model,want", [("syn_uni32k", 0b11111), ("botchan_uni8k", 0b11111), ("botchan_bpe4k", 0b11111),
                                        ("ref_ja_8k", None), ("wagahai_uni", None), ("ref_en_1k", None)])
def test_loader_filters(hostcore, make_hostcore, model, want):
    """What the loader derives from the charsmap for the byte-parallel paths: printable ASCII passes untouched
    (ascii_plain), TAB / LF / FF / CR normalize to one space (ws_ctl, nmt_nfkc models), and the per-character rule
    filter for 3-byte characters (cm_first3) never hides a rule -- every character it clears comes back from
    NormalizePrefix as itself, whatever follows it."""
    hc = make_hostcore(model_bytes(model))
    hostcore.hc_model_flags.argtypes = [ctypes.c_void_p]
    hostcore.hc_check_filters.argtypes = [ctypes.c_void_p]
    hostcore.hc_check_filters.restype = ctypes.c_int64
    flags = hostcore.hc_model_flags(hc.h)
    if want is not None:
        assert flags == want, bin(flags)
    assert hostcore.hc_check_filters(hc.h) == 0
    print(model, "flags", bin(flags))


def test_loader_rejects_bad_models(hostcore):
    import ctypes
    for blob in (b"", b"garbage\xff\xff", b"\x0a\x03abc"):
        code = ctypes.c_int(0)
        h = hostcore.hc_load(blob, len(blob), ctypes.byref(code))
        assert not h and code.value != 0
