"""The steps either side of EncodeAsIds: pieces with their spans
(SentencePieceText), batch Decode, dense / sparse layouts.

CPU part: the oracle restatement (oracle/spm_oracle.cc: encode_spt, decode_ids) and the
kernels' host-compiled arithmetic (tests/host_core.cc) against committed outputs of the
unmodified reference (tests/golden/ref_pieces.npz, generator make_pieces_fixtures.py) and the
reference's own DecodeTest vector (src/sentencepiece_processor_test.cc:466-541).
GPU part (-m gpu): the CUDA path through the C ABI against the same fixtures and against the
checker on larger seeded inputs.  Bar: byte-identical serialized protos / decoded text."""
import os
import sys

import numpy as np
import pytest

from conftest import EDGE_SENTENCES, GOLDEN, checker, model_bytes
from oracle import port

sys.path.insert(0, GOLDEN)
import make_fixtures as mf  # noqa: E402
import make_pieces_fixtures as mpf  # noqa: E402
import protowire as pw  # noqa: E402


@pytest.fixture(scope="module")
def ref_pieces():
    return np.load(os.path.join(GOLDEN, "ref_pieces.npz"))


@pytest.fixture(scope="module")
def small_corpora():
    return mf.corpora()


@pytest.fixture(scope="module")
def spt_mod():
    import spm_b200_pkg
    return spm_b200_pkg.load().spt      # pure-python formatter; importing needs no GPU


def protos_of(spt_mod, text, off, res, pieces):
    id_to_piece = lambda i: pieces[i]
    return [spt_mod.serialize(text[off[i]:off[i + 1]].tobytes(), spt_mod.sentence_pieces(i, text, off, res, id_to_piece))
            for i in range(len(off) - 1)]


def assert_protos(got, ref_pieces, name, text, off):
    proto, poff = ref_pieces[name + "__proto"], ref_pieces[name + "__poff"]
    assert len(got) == len(poff) - 1
    for i, g in enumerate(got):
        want = proto[poff[i]:poff[i + 1]].tobytes()
        if not want:       # the reference returned an error for this sentence: no pieces at all
            assert g == b"\n" + bytes([len(text[off[i]:off[i + 1]])]) + text[off[i]:off[i + 1]].tobytes() or \
                len(text[off[i]:off[i + 1]]) > 127, (name, i)
            continue
        assert g == want, (name, i, bytes(text[off[i]:off[i + 1]])[:80], g[:200], want[:200])


def model_pieces(mb):
    return [p[0] for p in pw.pieces_of(mb)]


CASE_IDS = [mpf.case_name(m, c, o) for m, c, _, o in mpf.CASES]


@pytest.mark.parametrize("model,corpus,first,opts", mpf.CASES, ids=CASE_IDS)
def test_oracle_pieces_match_reference(ref_pieces, small_corpora, spt_mod, model, corpus, first, opts):
    text, off = mpf.case_input(small_corpora, corpus, first)
    mb = model_bytes(model)
    o = port.OracleProcessor(mb)
    o.set_encode_extra_options(opts)
    res = o.encode_pieces_batch(text, off)
    assert_protos(protos_of(spt_mod, text, off, res, model_pieces(mb)), ref_pieces, mpf.case_name(model, corpus, opts),
                  text, off)


@pytest.mark.parametrize("model,corpus,first,opts", mpf.CASES, ids=CASE_IDS)
def test_core_pieces_match_reference(make_hostcore, ref_pieces, small_corpora, spt_mod, model, corpus, first, opts):
    """spm_core.cuh (aligned normalizer, piece ends, final_piece_at) compiled for the CPU."""
    text, off = mpf.case_input(small_corpora, corpus, first)
    mb = model_bytes(model)
    hc = make_hostcore(mb)
    res = hc.encode_pieces_batch(text, off, opts=port.pack_options(opts))
    # (inside: spans_from_ids, the fast half, must agree with the exact tier wherever it claims success)
    n_unknown_runs = 0 if corpus == "edge" or "wagahai" in (corpus, model.split("_")[0]) or model == "ref_ja_8k" else 1
    assert res["flagged"] <= (len(off) - 1) * (0.02 if n_unknown_runs else 1.0), res["flagged"]
    print(model, corpus, "flagged", res["flagged"], "plain", res["plain"], "of", len(off) - 1)
    # (inside too: the plain-sentence path -- length, text, spans -- against the exact tier on every sentence it takes)
    if corpus == "syn_en" and model.startswith("syn_"):
        assert res["plain"] >= 0.9 * (len(off) - 1), res["plain"]
    assert_protos(protos_of(spt_mod, text, off, res, model_pieces(mb)), ref_pieces, mpf.case_name(model, corpus, opts),
                  text, off)


@pytest.mark.parametrize("model,corpus,first,opts", mpf.CASES, ids=CASE_IDS)
def test_oracle_decode_matches_reference(ref_pieces, small_corpora, model, corpus, first, opts):
    name = mpf.case_name(model, corpus, opts)
    if name + "__dtext" not in ref_pieces.files:
        pytest.skip("the reference returned an error for a sentence of this case")
    text, off = mpf.case_input(small_corpora, corpus, first)
    o = port.OracleProcessor(model_bytes(model))
    o.set_encode_extra_options(opts)
    ids, io = o.encode_batch(text, off)
    dt, do = o.decode_batch(ids, io)
    assert np.array_equal(do, ref_pieces[name + "__doff"])
    assert np.array_equal(dt, ref_pieces[name + "__dtext"])


WS = "▁"


def ambiguous_model(model_type):
    """"QQ" and "QQQ" are pieces, "Q" alone is not: in "QQQQQ..." a run of unknown characters can
    end in several places, which only the Viterbi tie-break / the merge order settles -- the
    sentences spans_from_ids must hand back to the exact tier."""
    return pw.model(pw.base_pieces() + [(WS, -2.0), ("QQ", -1.5), ("QQQ", -3.5), ("a", -2.5), ("b", -2.5),
                                        ("ab", -3.0)], model_type=model_type)


AMBIGUOUS = [b"QQQQQ", b"QQQ", b"Q", b"QQ", b"aQQQQQb", b"QQQQQQQ ab QQQ", b"xQQQyQQQQz", b"QxQ", b"abQQQQQQQQQQQab",
             "日QQQ本QQQQQ".encode(), b"Q Q Q QQ QQQ QQQQ QQQQQ"]


@pytest.mark.parametrize("model_type", [pw.UNIGRAM, pw.BPE])
def test_core_pieces_ambiguous_unknown_runs(make_hostcore, model_type):
    from tools import synth
    mb = ambiguous_model(model_type)
    text, off = synth.pack_lines(AMBIGUOUS)
    hc = make_hostcore(mb)
    res = hc.encode_pieces_batch(text, off)          # (inside: fast spans == exact tier wherever they claim success)
    want = port.OracleProcessor(mb).encode_pieces_batch(text, off)
    for k in ("id_off", "norm_off", "ids", "begin", "end", "nbegin", "nend", "norm"):
        assert np.array_equal(res[k], want[k]), k
    if model_type == pw.UNIGRAM:
        assert res["flagged"] >= 1      # "QQQQQ": unk(Q) QQ QQ / QQ unk(Q) QQ / ... all fit the ids' shape


def decode_test_model():
    """DecodeMockModel of src/sentencepiece_processor_test.cc:467-494 as a real ModelProto."""
    return pw.model(pw.base_pieces() + [(WS + "ABC", -1.0), (WS + "DE", -1.0), ("F", -1.0), ("G" + WS + "H", -1.0)])


def test_oracle_decode_golden_vector():
    # sentencepiece_processor_test.cc:496-541 without the out-of-vocabulary piece "I"
    o = port.OracleProcessor(decode_test_model())
    ids = np.array([1, 3, 0, 4, 5, 6, 2], dtype=np.int32)
    t, to = o.decode_batch(ids, np.array([0, len(ids)]))
    assert t.tobytes() == b"ABC \xE2\x81\x87  DEFG H"
    # leading whitespace pieces are consumed only while the text is empty (:716-721)
    ids = np.array([1, 4, 4, 2, 3], dtype=np.int32)
    t, to = o.decode_batch(ids, np.array([0, 1, 5, 5]))
    assert t.tobytes() == b"DE DE ABC" and to.tolist() == [0, 0, 9, 9]


# ------------------------------------------------------------------ GPU: through the C ABI
gpu = pytest.mark.gpu


@gpu
@pytest.mark.parametrize("model,corpus,first,opts", mpf.CASES, ids=CASE_IDS)
def test_gpu_pieces_match_reference_fixture(spm_b200, ref_pieces, small_corpora, model, corpus, first, opts):
    text, off = mpf.case_input(small_corpora, corpus, first)
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(model_bytes(model))
    sp.SetEncodeExtraOptions(opts)
    try:
        got = sp.encode_as_serialized_protos(text, off)
    except spm_b200.StatusError as e:       # the reference's INTERNAL error on some sentence: arrays are still complete
        assert e.code == 13
        pytest.skip("case holds a sentence the reference rejects")
    assert_protos(got, ref_pieces, mpf.case_name(model, corpus, opts), text, off)


def assert_same_arrays(got, want):
    for k in ("id_off", "norm_off", "ids", "begin", "end", "nbegin", "nend", "norm"):
        assert np.array_equal(got[k], want[k]), k


@gpu
@pytest.mark.parametrize("model,corpus", [("botchan_uni8k", "botchan"), ("botchan_bpe4k", "botchan"),
                                          ("botchan_uni2k_ud", "botchan"), ("botchan_bpe2k_ud", "botchan"),
                                          ("syn_uni32k", "syn_en"), ("syn_uni32k", "syn_dirty"),
                                          ("syn_bpe32k", "syn_dirty"), ("ref_ja_8k", "wagahai"),
                                          ("wagahai_uni", "wagahai"), ("syn_uni32k", "wagahai")])
def test_gpu_pieces_match_oracle_arrays(spm_b200, corpora, model, corpus):
    """Whole corpora: every array of the C ABI against the CPU restatement (itself pinned above)."""
    text, off = corpora[corpus]
    if corpus.startswith("syn"):
        text, off = text[: off[6000]], off[:6001]
    mb = model_bytes(model)
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    o = port.OracleProcessor(mb)
    for opts in ("", "bos:eos", "reverse"):
        sp.SetEncodeExtraOptions(opts)
        o.set_encode_extra_options(opts)
        assert_same_arrays(sp.encode_pieces_packed(text, off), o.encode_pieces_batch(text, off))
    # the ids of the pieces path are the ids of the id path
    ids, io = sp.encode_packed(text, off)
    res = sp.encode_pieces_packed(text, off)
    assert np.array_equal(ids, res["ids"]) and np.array_equal(io, res["id_off"])


@gpu
@pytest.mark.parametrize("model_type", [pw.UNIGRAM, pw.BPE])
def test_gpu_pieces_ambiguous_unknown_runs(spm_b200, model_type):
    from tools import synth
    mb = ambiguous_model(model_type)
    text, off = synth.pack_lines(AMBIGUOUS * 40)
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    assert_same_arrays(sp.encode_pieces_packed(text, off), port.OracleProcessor(mb).encode_pieces_batch(text, off))
    if model_type == pw.UNIGRAM:
        assert sp.last_stats()["fallback_sentences"] >= 40     # the exact tier settled the ambiguous ones


@gpu
def test_gpu_pieces_single_sentence_api(spm_b200):
    mb = model_bytes("botchan_uni8k")
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    chk, kind = checker(mb)
    assert sp.EncodeAsPieces("") == []
    for s in EDGE_SENTENCES[:48]:
        text = np.frombuffer(s, dtype=np.uint8)
        off = np.array([0, len(s)], dtype=np.int64)
        if kind == "reference":
            want, _ = chk.encode_proto_batch(text, off)
            assert sp.EncodeAsSerializedProto(s) == want.tobytes(), s
    assert sp.EncodeAsPieces("Hello world.")[0].startswith(WS)


@gpu
def test_gpu_pieces_device_memory_and_capacity(spm_b200, corpora):
    import torch
    text, off = corpora["botchan"]
    mb = model_bytes("botchan_uni8k")
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    want = port.OracleProcessor(mb).encode_pieces_batch(text, off)
    dev = torch.device("cuda", 0)
    d_text = torch.from_numpy(text.copy()).to(dev)
    d_off = torch.from_numpy(off.copy()).to(dev)
    n = len(off) - 1
    cap, ncap = len(want["ids"]), len(want["norm"])
    arrs = [torch.empty(cap, dtype=torch.int32, device=dev) for _ in range(5)]
    id_off = torch.empty(n + 1, dtype=torch.int64, device=dev)
    norm = torch.empty(ncap, dtype=torch.uint8, device=dev)
    norm_off = torch.empty(n + 1, dtype=torch.int64, device=dev)
    sp.encode_pieces_device(d_text, d_off, *arrs, id_off, norm, norm_off)
    got = dict(zip(("ids", "begin", "end", "nbegin", "nend"), [a.cpu().numpy() for a in arrs]))
    got.update(id_off=id_off.cpu().numpy(), norm=norm.cpu().numpy(), norm_off=norm_off.cpu().numpy())
    assert_same_arrays(got, want)
    # too small: RESOURCE_EXHAUSTED, sizes to retry with in the offsets
    small = [torch.empty(cap - 5, dtype=torch.int32, device=dev) for _ in range(5)]
    with pytest.raises(spm_b200.StatusError) as e:
        sp.encode_pieces_device(d_text, d_off, *small, id_off, norm, norm_off)
    assert e.value.code == 8 and int(id_off[-1]) == cap
    with pytest.raises(spm_b200.StatusError) as e:
        sp.encode_pieces_device(d_text, d_off, *arrs, id_off, norm[: ncap - 1], norm_off)
    assert e.value.code == 8 and int(norm_off[-1]) == ncap
    # empty batch
    z = sp.encode_pieces_packed(np.zeros(0, dtype=np.uint8), np.zeros(1, dtype=np.int64))
    assert z["id_off"].tolist() == [0] and z["norm_off"].tolist() == [0] and len(z["ids"]) == 0


@gpu
@pytest.mark.parametrize("model,corpus,first,opts", mpf.CASES, ids=CASE_IDS)
def test_gpu_decode_matches_reference_fixture(spm_b200, ref_pieces, small_corpora, model, corpus, first, opts):
    name = mpf.case_name(model, corpus, opts)
    if name + "__dtext" not in ref_pieces.files:
        pytest.skip("the reference returned an error for a sentence of this case")
    text, off = mpf.case_input(small_corpora, corpus, first)
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(model_bytes(model))
    sp.SetEncodeExtraOptions(opts)
    ids, io = sp.encode_packed(text, off)
    dt, do = sp.decode_packed(ids, io)
    assert np.array_equal(do, ref_pieces[name + "__doff"])
    assert np.array_equal(dt, ref_pieces[name + "__dtext"])


@gpu
def test_gpu_decode_golden_vector_and_errors(spm_b200):
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(decode_test_model())
    assert sp.DecodeIds([1, 3, 0, 4, 5, 6, 2]).encode() == b"ABC \xE2\x81\x87  DEFG H"
    t, to = sp.decode_packed(np.array([1, 4, 4, 2, 3], dtype=np.int32), np.array([0, 1, 5, 5]))
    assert t.tobytes() == b"DE DE ABC" and to.tolist() == [0, 0, 9, 9]
    assert sp.DecodeIds([]) == ""
    t, to = sp.decode_packed(np.zeros(0, dtype=np.int32), np.zeros(1, dtype=np.int64))
    assert len(t) == 0 and to.tolist() == [0]
    with pytest.raises(spm_b200.StatusError) as e:
        sp.DecodeIds([1, 7])
    assert e.value.code == 11
    with pytest.raises(spm_b200.StatusError) as e:
        sp.DecodeIds([-1])
    assert e.value.code == 11


@gpu
@pytest.mark.parametrize("model,corpus", [("syn_uni32k", "syn_en"), ("syn_bpe32k", "syn_dirty"), ("ref_ja_8k", "wagahai"),
                                          ("botchan_uni2k_ud", "botchan")])
def test_gpu_decode_matches_checker_and_round_trip(spm_b200, corpora, model, corpus):
    """Whole corpora against the checker; plus the size-independent property: decoding the ids of
    a sentence without unknown pieces gives back its normalized text with U+2581 -> ' ' minus
    the dummy prefix."""
    text, off = corpora[corpus]
    mb = model_bytes(model)
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    ids, io = sp.encode_packed(text, off)
    dt, do = sp.decode_packed(ids, io)
    chk, kind = checker(mb)
    wt, wo = chk.decode_batch(ids, io)
    assert np.array_equal(do, wo) and np.array_equal(dt, wt)
    res = sp.encode_pieces_packed(text, off)
    unk = sp.unk_id()
    checked = 0
    for i in range(0, len(off) - 1, 7):
        sl = ids[io[i]:io[i + 1]]
        if (sl == unk).any() or len(sl) == 0:
            continue
        norm = res["norm"][res["norm_off"][i]:res["norm_off"][i + 1]].tobytes()
        want = norm.replace(WS.encode(), b" ")
        if norm.startswith(WS.encode()):
            want = want[1:]
        assert dt[do[i]:do[i + 1]].tobytes() == want
        checked += 1
    assert checked > 50 or corpus == "wagahai"      # nearly every wagahai line holds an unknown character


@gpu
def test_gpu_dense_and_sparse_layouts(spm_b200, corpora):
    """tensorflow/sentencepiece_processor_ops.cc:330-428 restated with numpy as the expectation."""
    import torch
    text, off = corpora["botchan"]
    sents = [text[off[i]:off[i + 1]].tobytes() for i in range(0, 700)]
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(model_bytes("botchan_uni8k"))
    sp.SetEncodeExtraOptions("reverse:bos:eos")      # RewritePieces order of the TF ops (:320-326)
    ids, io = sp.encode_batch(sents)
    lens = np.diff(io)
    mx = int(lens.max())
    want = np.full((len(sents), mx), 7, dtype=np.int32)
    for i in range(len(sents)):
        want[i, : lens[i]] = ids[io[i]:io[i + 1]]
    values, seq_len = sp.encode_dense(sents, pad=7)
    assert values.shape == (len(sents), mx)
    assert np.array_equal(values.cpu().numpy(), want) and np.array_equal(seq_len.cpu().numpy(), lens.astype(np.int32))
    dev = torch.device("cuda", 0)
    indices, vals, shape = sp.csr_to_sparse(torch.from_numpy(ids.copy()).to(dev), torch.from_numpy(io.copy()).to(dev))
    rows = np.repeat(np.arange(len(sents)), lens)
    cols = np.arange(len(ids)) - np.repeat(io[:-1], lens)
    assert shape == (len(sents), mx)
    assert np.array_equal(indices.cpu().numpy(), np.stack([rows, cols], axis=1))
    assert np.array_equal(vals.cpu().numpy(), ids)
    # empty rows only
    v, s = sp.csr_to_dense(torch.zeros(0, dtype=torch.int32, device=dev), torch.zeros(4, dtype=torch.int64, device=dev))
    assert v.shape == (3, 0) and s.cpu().tolist() == [0, 0, 0]
