"""E-step of the unigram trainer: the kernel's per-word function
against the reference's own closed-form test and against the compiled reference
(Lattice::PopulateMarginal over a TrainerModel).  Tolerance: 1e-5 relative (north star)."""
import ctypes
import math
import re

import numpy as np
import pytest

from conftest import model_bytes
import protowire as pw

REL = 1e-5


def pack_words(words):
    off = np.zeros(len(words) + 1, dtype=np.int64)
    np.cumsum([len(w) for w in words], out=off[1:])
    return np.frombuffer(b"".join(words), dtype=np.uint8).copy(), off


def hc_estep(L, pieces, scores, words, freq, all_freq=None):
    text, off = pack_words(words)
    scores = np.ascontiguousarray(scores, dtype=np.float32)
    freq = np.ascontiguousarray(freq, dtype=np.int64)
    expected = np.zeros(len(pieces), dtype=np.float64)
    out = np.zeros(2, dtype=np.float64)
    logz = np.zeros(len(words), dtype=np.float32)
    blob = b"".join(p + b"\0" for p in pieces)
    rc = L.hc_estep(blob, scores.ctypes.data, len(pieces), text.ctypes.data, off.ctypes.data, freq.ctypes.data,
                    len(words), float(all_freq or freq.sum()), expected.ctypes.data, out.ctypes.data,
                    logz.ctypes.data)
    assert rc == 0
    return expected, logz, out[0], int(round(out[1]))


def trainer_like_case(n_words=4000):
    """(pieces, scores, words, freq): the pieces of the botchan unigram fixture as a TrainerModel,
    the whitespace-split words of botchan with their counts (what SplitSentencesByWhitespace
    hands to RunEStep, src/trainer_interface.cc:465-477)."""
    from tools import synth
    ps = [(p, s) for p, s, t in pw.pieces_of(model_bytes("botchan_uni8k")) if t == pw.NORMAL]
    pieces = [p for p, _ in ps]
    scores = np.array([s for _, s in ps], dtype=np.float32)
    text = b"\n".join(synth.read_gz_lines("corpus_en_botchan.txt.gz")).decode("utf-8", "replace")
    cnt = {}
    for w in re.findall(r"\S+", text):
        w = ("▁" + w).encode("utf-8")
        cnt[w] = cnt.get(w, 0) + 1
    items = sorted(cnt.items(), key=lambda kv: (-kv[1], kv[0]))[:n_words]
    items.append(("▁".encode() + "日本語のテキスト".encode() * 12, 3))   # longer than the local tables
    items.append((b"\xe2\x96\x81zzzzqqq\xe2\x96\x85x", 2))                     # UNK nodes (U+2585 is not a piece)
    return pieces, scores, [w for w, _ in items], np.array([c for _, c in items], dtype=np.int64)


def check_against_reference(exp, logz, obj, nt, freq, rexp, rlogz, robj, rnt, rexp64):
    """North-star tolerance, SURVEY.md H9: per-word logZ and the expected counts against the
    float64 re-accumulation of the reference's OWN posteriors (oracle/ref_shim.cc ref_estep64)
    at 1e-5 relative; the Viterbi token count exactly.  The reference's float accumulator
    (expected[id] += ... in float over the whole corpus, src/unigram_model.cc:321) is itself
    further from that re-accumulation than 1e-5 on large counts: its delta is bounded loosely
    and printed, not held to the tolerance."""
    assert nt == rnt
    # the shim records what PopulateMarginal returns: freq * logZ (src/unigram_model.cc:327)
    f32 = freq.astype(np.float32)
    assert np.allclose(logz * f32, rlogz, rtol=REL, atol=0)
    floor = 1e-6                      # counts below this are sums of posteriors < 1e-6: absolute bound
    big = rexp64 > floor
    assert big.sum() > 1000
    rel = np.abs(exp[big] - rexp64[big]) / rexp64[big]
    assert rel.max() <= REL, "expected counts: max relative error %.3g vs the f64 re-accumulation" % rel.max()
    assert np.abs(exp[~big] - rexp64[~big]).max() <= floor * REL * 10 if (~big).any() else True
    # obj: the reference adds -Z / all_freq in float word by word; ours in double
    assert abs(obj - robj) <= 5e-5 * abs(robj)
    ref_rel = np.abs(rexp[big].astype(np.float64) - rexp64[big]) / rexp64[big]
    print("reference float accumulator vs its own f64 re-accumulation: max rel %.3g; ours: %.3g" % (
        ref_rel.max(), rel.max()))
    assert np.allclose(exp, rexp, rtol=1e-3, atol=1e-5)


def test_lse_table_accuracy(hostcore):
    """estep_softplus() (polynomial table) against log1pl(expl(d)) on a fine grid of [-50, 0]: the float that
    vmax + softplus rounds to then differs from the libm formula about once in 10^7 LogSumExp calls."""
    assert hostcore.hc_lse_table_error(2_000_003) < 1e-14


def test_closed_form_of_the_reference_test(hostcore):
    """src/unigram_model_test.cc:271-313 (LatticeTest.PopulateMarginalTest), tolerance 1e-3 there."""
    pieces = [b"A", b"B", b"C", b"AB", b"BC", b"ABC"]
    scores = [1.0, 1.2, 2.5, 3.0, 4.0, 2.0]
    exp, logz, obj, nt = hc_estep(hostcore, pieces, scores, [b"ABC"], [1])
    p1, p2, p3, p4 = math.exp(1.0 + 1.2 + 2.5), math.exp(3.0 + 2.5), math.exp(1.0 + 4.0), math.exp(2.0)
    Z = p1 + p2 + p3 + p4
    want = [(p1 + p3) / Z, p1 / Z, (p1 + p2) / Z, p2 / Z, p3 / Z, p4 / Z]
    assert np.allclose(exp, want, atol=1e-5)
    assert abs(logz[0] - math.log(Z)) < 1e-5
    assert nt == 2          # Viterbi: AB C (5.5) beats A BC (5.0), A B C (4.7), ABC (2.0)
    assert abs(obj + math.log(Z)) < 1e-5


def test_hostcore_matches_reference_estep(hostcore):
    from oracle import ref
    if not ref.available() or not hasattr(ref.lib(), "ref_estep"):
        pytest.skip("reference library with the trainer is not built")
    pieces, scores, words, freq = trainer_like_case()
    rexp, rlogz, robj, rnt, rexp64 = ref.estep64(pieces, scores, words, freq)
    exp, logz, obj, nt = hc_estep(hostcore, pieces, scores, words, freq)
    check_against_reference(exp, logz, obj, nt, freq, rexp, rlogz, robj, rnt, rexp64)


@pytest.mark.gpu
def test_gpu_estep_matches_reference(spm_b200):
    from oracle import ref
    if not hasattr(ref.lib(), "ref_estep"):
        pytest.skip("reference library was built without the trainer")
    pieces, scores, words, freq = trainer_like_case()
    rexp, rlogz, robj, rnt, rexp64 = ref.estep64(pieces, scores, words, freq)
    es = spm_b200.UnigramEStep(pieces, scores)
    text, off = pack_words(words)
    exp, obj, nt, logz = es.run(text, off, freq, want_logz=True)
    check_against_reference(exp, logz, obj, nt, freq, rexp, rlogz, robj, rnt, rexp64)
    # two shards summed = the whole (what the all-reduce does across GPUs)
    from importlib import import_module
    sh = import_module("sentencepiece-comments_b200.sharding")
    tot = np.zeros_like(exp)
    o2, n2 = 0.0, 0
    for lo, hi in sh.estep_shard_ranges(off, freq, 2):
        e, o, n = es.run(text[off[lo]:off[hi]], off[lo:hi + 1] - off[lo], freq[lo:hi], all_freq=float(freq.sum()))
        tot += e
        o2 += o
        n2 += n
    assert n2 == nt and np.allclose(tot, exp, rtol=1e-9) and abs(o2 - obj) < 1e-9 * abs(obj)
