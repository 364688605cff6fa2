"""The compiled C++ binding (include/spm_cuda_processor.h): sentencepiece::CudaSentencePieceProcessor
built against the UNMODIFIED reference header and called through `SentencePieceProcessor *`
(tests/facade_driver.cc), with the reference base class as the checker -- plus the load-time
self-test of ModelProto.self_test_data (src/sentencepiece_processor.cc:134-154)."""
import ctypes
import os

import numpy as np
import pytest

from conftest import EDGE_SENTENCES, ROOT, model_bytes
import build_facade


@pytest.fixture(scope="module")
def facade():
    path = build_facade.build()
    if not path or not os.path.exists(path):
        pytest.skip("tests/_build/libfacade_test.so is not built (needs the reference sources)")
    L = ctypes.CDLL(path)
    L.facade_run.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                             ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_int]
    return L


def run(L, mb, lines, opts=b"", vocab=None):
    from tools import synth
    text, off = synth.pack_lines(lines)
    text = np.concatenate([text, np.zeros(8, dtype=np.uint8)])
    err = ctypes.create_string_buffer(512)
    vb = b"".join(v + b"\0" for v in vocab) if vocab is not None else None
    rc = L.facade_run(mb, len(mb), text.ctypes.data, off.ctypes.data, len(lines), opts, vb,
                      len(vocab) if vocab is not None else -1, err, 512)
    return rc, err.value.decode("utf-8", "replace")


def test_facade_compiles_and_fails_loudly_without_a_gpu(facade):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    rc, msg = run(facade, model_bytes("botchan_uni8k"), [b"hello world"])
    assert rc == -1 and "no CUDA device" in msg       # the subclass has no CPU path either


@pytest.mark.gpu
@pytest.mark.parametrize("model,corpus,opts", [
    ("botchan_uni8k", "botchan", b""), ("botchan_bpe4k", "botchan", b""), ("botchan_uni2k_ud", "botchan", b"bos:eos"),
    ("ref_ja_8k", "wagahai", b"reverse"), ("syn_uni32k", "syn_dirty", b"eos:reverse:bos"), ("ref_en_1k", "botchan", b""),
    ("botchan_uni8k", "botchan", b"no_such_option"),
])
def test_facade_matches_base_class(facade, model, corpus, opts):
    from tools import synth
    lines = {"botchan": synth.read_gz_lines("corpus_en_botchan.txt.gz")[:1200],
             "wagahai": synth.read_gz_lines("corpus_ja_wagahai.txt.gz")[:300]}.get(corpus)
    if lines is None:
        t, o = synth.EnglishCorpus(dirty=True).generate(1500, first=9)
        lines = [bytes(t[o[i]:o[i + 1]]) for i in range(1500)]
    lines = lines + [s for s in EDGE_SENTENCES if b"\0" not in s]     # (the driver hands sentences over as std::string)
    rc, msg = run(facade, model_bytes(model), lines, opts)
    assert rc == 0, msg


@pytest.mark.gpu
@pytest.mark.parametrize("model", ["botchan_uni8k", "botchan_bpe4k"])
def test_facade_set_vocabulary(facade, spm_b200, model):
    from tools import synth
    mb = model_bytes(model)
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    vocab = [sp._piece_bytes(i) for i in range(sp.GetPieceSize()) if i % 3 != 0]
    vocab = [v for v in vocab if b"\0" not in v]
    rc, msg = run(facade, mb, synth.read_gz_lines("corpus_en_botchan.txt.gz")[:800], b"", vocab)
    assert rc == 0, msg


def with_self_test(mb, samples):
    """mb + ModelProto.self_test_data (field 4: repeated Sample {input = 1, expected = 2})."""
    import protowire as pw
    body = b"".join(pw._ld(1, pw._ld(1, i) + pw._ld(2, e)) for i, e in samples)
    return mb + pw._ld(4, body)


@pytest.mark.gpu
@pytest.mark.parametrize("model", ["botchan_uni8k", "botchan_bpe4k", "ref_ja_8k"])
def test_self_test_data_at_load(spm_b200, model):
    from oracle import ref
    mb = model_bytes(model)
    sp = spm_b200.SentencePieceProcessor()
    sp.LoadFromSerializedProto(mb)
    inputs = ["Hello world.", " I  saw a girl with\ta telescope. ", "吾輩は猫である。名前はまだ無い。", "ｸﾞｰｸﾞﾙ ①②③", ""]
    good = [(s.encode(), " ".join(sp.EncodeAsPieces(s)).encode()) for s in inputs]
    ok = with_self_test(mb, good)
    sp2 = spm_b200.SentencePieceProcessor()
    sp2.LoadFromSerializedProto(ok)                       # every sample reproduces: loads
    assert sp2.EncodeAsIds(inputs[0]) == sp.EncodeAsIds(inputs[0])
    bad = with_self_test(mb, good[:2] + [(good[2][0], good[2][1] + b" x")] + good[3:])
    with pytest.raises(spm_b200.StatusError) as e:
        spm_b200.SentencePieceProcessor().LoadFromSerializedProto(bad)
    assert e.value.code == 13 and "Self-test failures" in str(e.value) and "1/5" in str(e.value)
    if ref.available():                                   # the reference takes the same two decisions
        ref.RefProcessor(ok)
        with pytest.raises(RuntimeError):
            ref.RefProcessor(bad)
