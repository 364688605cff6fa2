"""GPU parity where round 1 had none (VERDICT r1 "What's weak" 1-3):

  * k_encode_fast itself producing the ids (lists larger than warp_route, and small lists with
    the routing switched off through the per-handle option);
  * configs 2 and 3 at 2 M sentences, HOST (pipelined chunks) and DEVICE (one chunk) calls;
  * the memo table forced small: probe exhaustion, more new words than the list holds;
  * every tier switch kept for measurements (memo off, warp kernel off, per-sentence scan,
    exact pieces tier) as a parity case of its own;
  * the multi-chunk HOST pipeline on small batches (chunk threshold lowered);
  * more near-tied Viterbi candidates than the exact tier's first capacity (retry).

The checker is the unmodified reference (oracle/_ref) when present, else the CPU restatement.
Bit-exact ids and offsets are required throughout."""
import os

import numpy as np
import pytest

from conftest import NEAR_TIE_SENTENCES, checker, model_bytes, near_tie_model
from test_gpu_parity import assert_same

pytestmark = pytest.mark.gpu


def ref_encode_all(chk, text, off):
    try:
        return chk.encode_batch(text, off, threads=len(os.sched_getaffinity(0)))
    except TypeError:
        return chk.encode_batch(text, off)


def load(spm_b200, mb, **opts):
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    for k, v in opts.items():
        sp.set_option(k, v)
    return sp


def inner_ws_model(name):
    """`name` + one NORMAL piece with U+2581 inside: lattice paths cross word boundaries, so the
    loader must refuse the word memo (memo_ok == false) and every sentence takes the Viterbi
    kernels.  (A serialized ModelProto may carry repeated fields in any order: appending one
    more `pieces` entry is a valid model whose new piece gets the next id.)"""
    import protowire as pw
    return model_bytes(name) + pw._ld(1, pw.piece("▁of▁the", -6.0))


# ---------------------------------------------------------------- k_encode_fast produces the ids
@pytest.mark.parametrize("model,corpus,opts", [
    ("ref_ja_8k", "wagahai", {"warp_route": 0}),
    ("wagahai_uni", "wagahai", {"warp_route": 0}),
    ("syn_uni32k", "syn_en", {"warp_route": 0, "memo": 0}),
    ("syn_uni32k", "syn_dirty", {"warp_route": 0, "memo": 0}),
    ("botchan_uni8k", "botchan", {"warp_route": 0, "memo": 0}),
    ("botchan_uni2k_ud", "botchan", {"warp_route": 0, "memo": 0}),
    ("syn_uni32k", "wagahai", {"warp_route": 0}),
    ("ref_ja_8k", "wagahai", {"warp_route": 0, "warp": 0}),      # flagged lines -> general tier, long lines -> k_encode_fast
    ("syn_uni32k", "syn_dirty", {"warp_route": 0, "warp": 0, "memo": 0}),
])
def test_encode_fast_kernel_small_lists(spm_b200, corpora, model, corpus, opts):
    mb = model_bytes(model)
    sp = load(spm_b200, mb, **opts)
    chk, _ = checker(mb)
    text, off = corpora[corpus]
    assert_same(sp.encode_packed(text, off), ref_encode_all(chk, text, off), text, off)
    st = sp.last_stats()
    assert st["viterbi_tier_sentences"] > 1000, st      # the thread-per-sentence kernel did the work


@pytest.mark.parametrize("model,kind", [("ref_ja_8k", "wagahai"), ("wagahai_uni", "wagahai"),
                                        ("inner_ws", "syn_en")])
def test_encode_fast_kernel_300k(spm_b200, model, kind):
    """More than 262,144 leftover sentences in ONE chunk with default options: k_encode_fast keeps the list."""
    from tools import synth
    n = 300_000
    if kind == "wagahai":
        text, off = synth.ResampledCorpus().generate(n, first=17)
        mb = model_bytes(model)
    else:
        text, off = synth.EnglishCorpus().generate(n, first=5_000_000)
        mb = inner_ws_model("syn_uni32k")
    sp = load(spm_b200, mb)
    chk, _ = checker(mb)
    import torch
    want = ref_encode_all(chk, text, off)
    dt, do = torch.from_numpy(text).cuda(), torch.from_numpy(off).cuda()
    ids = torch.empty(int(want[1][-1]) + 8, dtype=torch.int32, device="cuda")
    ido = torch.empty(len(off), dtype=torch.int64, device="cuda")
    sp.encode_device(dt, do, ids, ido)
    torch.cuda.synchronize()
    st = sp.last_stats()
    assert_same((ids[: int(want[1][-1])].cpu().numpy(), ido.cpu().numpy()), want, text, off)
    assert st["viterbi_tier_sentences"] > 262_144, st
    if kind != "wagahai":
        assert st["stage_ms"]["memo_collect"] == 0.0, "a model with U+2581 inside a piece must not use the word memo"
        # the same through the HOST call
        assert_same(sp.encode_packed(text, off), want, text, off)


# ---------------------------------------------------------------- configs 2 / 3 at 2 M sentences
@pytest.mark.parametrize("model", ["syn_uni32k", "syn_bpe32k"])
def test_config_2m_parity(spm_b200, model):
    """2 M sentences of the benched corpus: HOST call (12 / 5 pipelined chunks, carried CSR offsets,
    incremental copy-out) and DEVICE call (one chunk, the memo table at 2^22 slots) against the
    reference, every id."""
    import torch
    from tools import synth
    n = 2_000_000
    text, off = synth.EnglishCorpus().generate(n)
    mb = model_bytes(model)
    sp = load(spm_b200, mb)
    chk, _ = checker(mb)
    want = ref_encode_all(chk, text, off)
    total = int(want[1][-1])
    h_text, h_off = torch.from_numpy(text).pin_memory(), torch.from_numpy(off).pin_memory()
    h_ids = torch.empty(total + 64, dtype=torch.int32).pin_memory()
    h_ido = torch.empty(n + 1, dtype=torch.int64).pin_memory()
    sp.encode_pinned(h_text, h_off, h_ids, h_ido)
    assert_same((h_ids[:total].numpy(), h_ido.numpy()), want, text, off)
    dt, do = h_text.cuda(), h_off.cuda()
    ids = torch.empty(total + 8, dtype=torch.int32, device="cuda")
    ido = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    sp.encode_device(dt, do, ids, ido)
    torch.cuda.synchronize()
    st = sp.last_stats()
    assert_same((ids[:total].cpu().numpy(), ido.cpu().numpy()), want, text, off)
    assert st["fallback_sentences"] < n // 1000 and st["viterbi_tier_sentences"] < n // 50, st
    # exact capacity is enough; one id less is RESOURCE_EXHAUSTED with the size to retry with
    small = torch.empty(total - 1, dtype=torch.int32).pin_memory()
    with pytest.raises(spm_b200.StatusError) as e:
        sp.encode_pinned(h_text, h_off, small, h_ido)
    assert e.value.code == 8 and int(h_ido[n]) == total


@pytest.mark.parametrize("model", ["syn_uni32k", "syn_bpe32k"])
def test_dirty_1m_host_chunks(spm_b200, model):
    """2^20 + 1000 sentences with NFKC / whitespace / malformed noise through the chunked HOST call."""
    import torch
    from tools import synth
    n = (1 << 20) + 1000
    text, off = synth.EnglishCorpus(dirty=True).generate(n, first=123)
    mb = model_bytes(model)
    sp = load(spm_b200, mb)
    chk, _ = checker(mb)
    want = ref_encode_all(chk, text, off)
    assert_same(sp.encode_packed(text, off), want, text, off)


# ---------------------------------------------------------------- the memo table forced small
@pytest.mark.parametrize("model,corpus", [("syn_uni32k", "syn_en"), ("syn_bpe32k", "syn_en"),
                                          ("syn_uni32k", "syn_dirty"), ("botchan_uni8k", "botchan")])
@pytest.mark.parametrize("bits", [12, 14])
def test_small_memo_table(spm_b200, corpora, model, corpus, bits):
    """2^12 slots for ~40 k distinct words: probe sequences run out (24 probes), more words are
    claimed than new_list holds (n_new > max_new: claimed, never encoded, never served)."""
    mb = model_bytes(model)
    sp = load(spm_b200, mb, memo_bits=bits)
    chk, _ = checker(mb)
    text, off = corpora[corpus]
    want = ref_encode_all(chk, text, off)
    assert_same(sp.encode_packed(text, off), want, text, off)
    st = sp.last_stats()
    if corpus != "botchan":
        assert st["viterbi_tier_sentences"] > 1000, st       # the full table really turned sentences away
    sp.set_option("scan_per_sentence", 1)
    assert_same(sp.encode_packed(text, off), want, text, off)


def test_memo_overflow_rows(spm_b200):
    """Words of 6..24 ids (overflow rows of the memo results) and words at the key limits."""
    from tools import synth
    mb = model_bytes("botchan_uni8k")
    sp = load(spm_b200, mb)
    chk, _ = checker(mb)
    rng = np.random.default_rng(3)
    alpha = b"qxzjkvwQXZJ0123456789#@"
    lines = []
    for i in range(4000):
        ws = []
        for _ in range(int(rng.integers(1, 12))):
            n = int(rng.integers(1, 27))
            ws.append(bytes(alpha[int(k)] for k in rng.integers(0, len(alpha), size=n)))
        lines.append(b" ".join(ws))
    text, off = synth.pack_lines(lines)
    assert_same(sp.encode_packed(text, off), ref_encode_all(chk, text, off), text, off)


# ---------------------------------------------------------------- tier switches as parity cases
@pytest.mark.parametrize("model,corpus", [("syn_uni32k", "syn_en"), ("syn_uni32k", "syn_dirty"),
                                          ("syn_bpe32k", "syn_en"), ("syn_bpe32k", "syn_dirty"),
                                          ("ref_ja_8k", "wagahai"), ("botchan_uni2k_ud", "botchan"),
                                          ("botchan_bpe2k_ud", "botchan")])
@pytest.mark.parametrize("opts", [{"memo": 0}, {"warp": 0}, {"scan_per_sentence": 1}, {"memo": 0, "warp": 0},
                                  {"emit_blocks": 1}, {"max_chunk": 777}, {"fast_side": 0, "warp_route": 0},
                                  {"mid_lines": 1, "warp_route": 0}])
def test_option_paths(spm_b200, corpora, model, corpus, opts):
    mb = model_bytes(model)
    sp = load(spm_b200, mb, **opts)
    for k, v in opts.items():
        assert sp.get_option(k) == v
    chk, _ = checker(mb)
    text, off = corpora[corpus]
    assert_same(sp.encode_packed(text, off), ref_encode_all(chk, text, off), text, off)


def test_unknown_option(spm_b200):
    sp = load(spm_b200, model_bytes("botchan_uni8k"))
    with pytest.raises(spm_b200.StatusError) as e:
        sp.set_option("no_such_knob", 1)
    assert e.value.code == 5


@pytest.mark.parametrize("model,corpus", [("syn_uni32k", "syn_dirty"), ("botchan_bpe4k", "botchan"),
                                          ("ref_ja_8k", "wagahai"), ("botchan_uni2k_ud", "botchan")])
def test_pieces_exact_tier(spm_b200, corpora, model, corpus):
    """encode_batch_pieces with every sentence through the exact tier == the default path == the oracle."""
    from oracle import port
    mb = model_bytes(model)
    text, off = corpora[corpus]
    m = min(3000, len(off) - 1)
    text, off = text[: off[m]], off[: m + 1]
    want = port.OracleProcessor(mb).encode_pieces_batch(text, off)
    for opts in ({"pieces_exact": 0}, {"pieces_exact": 1}, {"pieces_plain": 0}):
        sp = load(spm_b200, mb, **opts)
        got = sp.encode_pieces_packed(text, off)
        for k in ("id_off", "norm_off", "ids", "begin", "end", "nbegin", "nend", "norm"):
            assert np.array_equal(got[k], want[k]), (k, opts)


@pytest.mark.parametrize("model,opts", [("syn_uni32k", 0), ("syn_bpe32k", 0), ("syn_uni32k", 1), ("botchan_uni8k", 0)])
def test_pieces_plain_sentences(spm_b200, model, opts):
    """The plain-sentence half of encode_batch_pieces (no normalizer, no norm_to_orig) on 200 k clean and noisy
    synthetic sentences, with and without extra options, against the path that normalizes every sentence."""
    from tools import synth
    n = 200_000
    ct, co = synth.EnglishCorpus().generate(n // 2, first=3_000_000)
    dt, do = synth.EnglishCorpus(dirty=True).generate(n // 2, first=11)
    text = np.concatenate([ct, dt])
    off = np.concatenate([co, do[1:] + co[-1]])
    mb = model_bytes(model)
    extra = ["", "bos:eos:reverse"][opts]
    res = []
    for plain in (1, 0):
        sp = load(spm_b200, mb, pieces_plain=plain)
        sp.SetEncodeExtraOptions(extra)
        res.append(sp.encode_pieces_packed(text, off))
    for k in ("id_off", "norm_off", "ids", "begin", "end", "nbegin", "nend", "norm"):
        assert np.array_equal(res[0][k], res[1][k]), k


# ---------------------------------------------------------------- multi-chunk HOST pipeline on small batches
@pytest.mark.parametrize("model,corpus", [("syn_uni32k", "syn_en"), ("syn_uni32k", "syn_dirty"),
                                          ("syn_bpe32k", "syn_dirty"), ("ref_ja_8k", "wagahai"),
                                          ("ref_en_1k", "botchan")])
@pytest.mark.parametrize("chunks", [0, 2, 64])
def test_host_chunk_pipeline(spm_b200, corpora, model, corpus, chunks):
    import torch
    mb = model_bytes(model)
    sp = load(spm_b200, mb, host_chunk_min=1, host_chunks=chunks)
    chk, _ = checker(mb)
    text, off = corpora[corpus]
    want = ref_encode_all(chk, text, off)
    assert_same(sp.encode_packed(text, off), want, text, off)
    # ... and it equals the single-chunk DEVICE result
    dt, do = torch.from_numpy(text).cuda(), torch.from_numpy(off).cuda()
    ids = torch.empty(int(want[1][-1]) + 8, dtype=torch.int32, device="cuda")
    ido = torch.empty(len(off), dtype=torch.int64, device="cuda")
    sp.encode_device(dt, do, ids, ido)
    torch.cuda.synchronize()
    assert_same((ids[: int(want[1][-1])].cpu().numpy(), ido.cpu().numpy()), want, text, off)
    # too small a capacity in the chunked call: RESOURCE_EXHAUSTED, offsets complete
    cap = int(want[1][-1]) // 2
    ido_h = np.zeros(len(off), dtype=np.int64)
    small = np.empty(cap, dtype=np.int32)
    with pytest.raises(spm_b200.StatusError) as e:
        sp.encode_pinned(text, off, small, ido_h)
    assert e.value.code == 8
    assert np.array_equal(ido_h, want[1])
    assert np.array_equal(small, want[0][:cap])


# ---------------------------------------------------------------- near-tied candidates beyond the first capacity
@pytest.mark.parametrize("opts", [{}, {"warp": 0}, {"memo": 0}, {"warp_route": 0}])
def test_near_tie_retry(spm_b200, opts):
    from tools import synth
    mb = near_tie_model()
    sp = load(spm_b200, mb, **opts)
    chk, _ = checker(mb)
    lines = NEAR_TIE_SENTENCES + [b"a" * int(k) for k in np.random.default_rng(1).integers(1, 400, size=500)]
    text, off = synth.pack_lines(lines)
    want = ref_encode_all(chk, text, off)
    assert_same(sp.encode_packed(text, off), want, text, off)
    assert sp.last_stats()["fallback_sentences"] >= 4     # they reached the exact tier (and its retry)
    # the pieces path takes the same retry
    from oracle import port
    w = port.OracleProcessor(mb).encode_pieces_batch(text, off)
    got = sp.encode_pieces_packed(text, off)
    for k in ("id_off", "ids", "begin", "end", "nbegin", "nend"):
        assert np.array_equal(got[k], w[k]), k


def test_release_workspace_and_small_batch_footprint(spm_b200, corpora):
    """A small batch with a 9 KB line must not size the per-thread back-pointer rows of
    k_encode_fast (3.76 GB in round 1); release_workspace gives the rest back."""
    mb = model_bytes("botchan_uni8k")
    sp = load(spm_b200, mb)
    text, off = corpora["edge"]
    want = sp.encode_packed(text, off)
    st = sp.last_stats()
    assert st["workspace_bytes"] < (1 << 30), st
    sp.release_workspace()
    got = sp.encode_packed(text, off)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])


def test_caller_device_is_restored(spm_b200):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    sp = spm_b200.SentencePieceProcessor(device=1)
    torch.cuda.set_device(0)
    sp.LoadFromSerializedProto(model_bytes("botchan_uni8k"))
    assert sp.EncodeAsIds("hello world")
    assert torch.cuda.current_device() == 0
    x = torch.ones(4, device="cuda")
    assert x.device.index == 0
