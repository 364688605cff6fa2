#!/usr/bin/env python3
"""Golden fixtures of the steps either side of EncodeAsIds, from the UNMODIFIED reference
(oracle/_ref/libspm_ref.so):

  tests/golden/ref_pieces.npz
     <case>__proto / __poff   EncodeAsSerializedProto of every sentence (back to back, CSR)
     <case>__dtext / __doff   Decode(ids) of the reference's own ids of every sentence

A case is model__corpus[__options].  Run in the build container only; the tests read the
fixture, never the reference tree.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref  # noqa: E402
import conftest  # noqa: E402
import make_fixtures as mf  # noqa: E402

CASES = [  # (model, corpus, first n sentences, extra options)
    ("botchan_uni8k", "botchan", 400, ""),
    ("botchan_uni8k", "botchan", 200, "bos:eos"),
    ("botchan_uni8k", "botchan", 200, "reverse:bos"),
    ("botchan_bpe4k", "botchan", 400, ""),
    ("botchan_bpe4k", "botchan", 200, "eos:reverse"),
    ("botchan_uni2k_ud", "botchan", 300, ""),
    ("botchan_bpe2k_ud", "botchan", 300, ""),
    ("ref_en_1k", "botchan", 200, ""),
    ("ref_ja_8k", "wagahai", 150, ""),
    ("wagahai_uni", "wagahai", 100, ""),
    ("syn_uni32k", "syn_dirty", 400, ""),
    ("syn_bpe32k", "syn_dirty", 400, ""),
    ("syn_uni32k", "wagahai", 60, ""),      # long unknown runs
]
CASES += [(m, "edge", None, "") for m in mf.EDGE_MODELS]
CASES += [("botchan_uni8k", "edge", None, "bos:eos:reverse")]


def case_name(model, corpus, opts):
    return "%s__%s%s" % (model, corpus, "__" + opts.replace(":", "_") if opts else "")


def case_input(cs, corpus, first):
    text, off = cs[corpus]
    if first is not None:
        text, off = text[: off[first]], off[: first + 1]
    return text, off


def main():
    cs = mf.corpora()
    out = {}
    for model, corpus, first, opts in CASES:
        text, off = case_input(cs, corpus, first)
        rp = ref.RefProcessor(conftest.model_bytes(model))
        if opts:
            rp.set_encode_extra_options(opts)
        proto, poff = rp.encode_proto_batch(text, off, threads=8)
        name = case_name(model, corpus, opts)
        out[name + "__proto"] = proto.copy()
        out[name + "__poff"] = poff
        try:
            ids, io = rp.encode_batch(text, off, threads=8)
        except RuntimeError:     # a sentence with the reference's INTERNAL error: ids of the others
            ids, io = None, None
        if ids is not None:
            dt, do = rp.decode_batch(ids, io, threads=8)
            out[name + "__dtext"] = dt.copy()
            out[name + "__doff"] = do
        print(name, len(off) - 1, "sentences", len(proto), "proto bytes")
    np.savez_compressed(os.path.join(HERE, "ref_pieces.npz"), **out)


if __name__ == "__main__":
    main()
