#!/usr/bin/env python3
"""Generates the committed golden fixtures from the UNMODIFIED reference
(oracle/_ref/libspm_ref.so, built by oracle/build_ref.py from /root/reference):

  tests/golden/charsmap_space_rules.bin  the charsmap of normalizer_test.cc:181-200
                                         (a->" A", b->"B", c->"D E", d->" F G "),
                                         compiled by the reference's Builder::CompileCharsMap
  tests/golden/ref_ids.npz               reference ids / CSR offsets on seeded inputs
  tests/golden/ref_norm.npz              reference normalized strings of the edge sentences

Run in the build container only (needs /root/reference for the first build of
oracle/_ref).  The tests read the fixtures, never the reference tree.
"""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref  # noqa: E402
from tools import synth  # noqa: E402
import conftest  # noqa: E402

CASES = [  # (model, corpus, first n sentences or None)
    ("botchan_uni8k", "botchan", None),   # config 1: 4,288 lines, 67,126 ids
    ("botchan_bpe4k", "botchan", None),
    ("botchan_uni2k_ud", "botchan", 1500),
    ("botchan_bpe2k_ud", "botchan", 1500),
    ("ref_en_1k", "botchan", 1500),
    ("ref_ja_8k", "wagahai", None),       # config 4 fixture model
    ("wagahai_uni", "wagahai", 800),
    ("syn_uni32k", "syn_en", 4000),       # config 2 sample
    ("syn_uni32k", "syn_dirty", 4000),
    ("syn_bpe32k", "syn_en", 4000),       # config 3 sample
    ("syn_bpe32k", "syn_dirty", 4000),
]
EDGE_MODELS = ["botchan_uni8k", "botchan_bpe4k", "botchan_uni2k_ud", "botchan_bpe2k_ud", "syn_uni32k",
               "syn_bpe32k", "wagahai_uni", "ref_ja_8k", "ref_en_1k"]


def corpora():
    return {"botchan": synth.pack_lines(synth.read_gz_lines("corpus_en_botchan.txt.gz")),
            "wagahai": synth.pack_lines(synth.read_gz_lines("corpus_ja_wagahai.txt.gz")),
            "syn_en": synth.EnglishCorpus().generate(4000),
            "syn_dirty": synth.EnglishCorpus(dirty=True).generate(4000, first=777),
            "edge": synth.pack_lines(conftest.EDGE_SENTENCES)}


def main():
    L = ref.lib()
    L.ref_compile_charsmap.restype = ctypes.c_int64
    L.ref_compile_charsmap.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int64]
    rules = b"a\0 A\0b\0B\0c\0D E\0d\0 F G \0"
    buf = ctypes.create_string_buffer(1 << 20)
    n = L.ref_compile_charsmap(rules, 4, buf, len(buf))
    assert n > 0
    with open(os.path.join(HERE, "charsmap_space_rules.bin"), "wb") as f:
        f.write(buf.raw[:n])

    cs = corpora()
    out = {}
    for model, corpus, first in CASES:
        text, off = cs[corpus]
        if first is not None:
            text, off = text[: off[first]], off[: first + 1]
        ids, io = ref.RefProcessor(conftest.model_bytes(model)).encode_batch(text, off, threads=8)
        out["%s__%s__ids" % (model, corpus)] = ids.astype(np.int32)
        out["%s__%s__off" % (model, corpus)] = io.astype(np.int64)
        print(model, corpus, len(off) - 1, "sentences", len(ids), "ids")
    text, off = cs["edge"]
    norm = {}
    for model in EDGE_MODELS:
        rp = ref.RefProcessor(conftest.model_bytes(model))
        ids, io = rp.encode_batch(text, off)
        out["%s__edge__ids" % model] = ids.astype(np.int32)
        out["%s__edge__off" % model] = io.astype(np.int64)
        strs = [rp.normalize(s)[0] for s in conftest.EDGE_SENTENCES]
        norm[model + "__text"] = np.frombuffer(b"".join(strs), dtype=np.uint8)
        norm[model + "__off"] = np.cumsum([0] + [len(s) for s in strs]).astype(np.int64)
    np.savez_compressed(os.path.join(HERE, "ref_ids.npz"), **out)
    np.savez_compressed(os.path.join(HERE, "ref_norm.npz"), **norm)


if __name__ == "__main__":
    main()
