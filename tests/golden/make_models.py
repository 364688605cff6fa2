#!/usr/bin/env python3
"""Trains the fixture models with the REFERENCE trainer (oracle/_ref, i.e. the
reference's own SentencePieceTrainer::Train compiled from /root/reference) and
writes them to tests/golden/models/.  Run once in the build container; the
.model files are committed (they are the inputs parity is defined on).

  botchan_uni8k   config 1: --vocab_size=8000 --model_type=unigram on botchan.txt
                  (--hard_vocab_limit=false: the corpus only supports ~5.2k
                  pieces, src/trainer_interface.cc:516-527)
  syn_uni32k      config 2/5: unigram 32k, nmt_nfkc, synthetic English-like corpus
  syn_bpe32k      config 3: BPE 32k on the same corpus
  wagahai_uni     config 4: unigram CJK on wagahaiwa_nekodearu.txt
  botchan_bpe4k   small BPE model for fast tests
  botchan_uni2k_ud  unigram with user-defined symbols + control symbols
"""
import gzip
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402
from tools import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "models")


def gunzip_to(name, path):
    with gzip.open(os.path.join(ROOT, "tests", "golden", "data", name), "rb") as f, open(path, "wb") as g:
        g.write(f.read())


def main(only=None):
    os.makedirs(OUT, exist_ok=True)
    with tempfile.TemporaryDirectory() as td:
        bot = os.path.join(td, "botchan.txt")
        wag = os.path.join(td, "wagahai.txt")
        gunzip_to("corpus_en_botchan.txt.gz", bot)
        gunzip_to("corpus_ja_wagahai.txt.gz", wag)
        syn = os.path.join(td, "syn.txt")
        text, off = synth.EnglishCorpus().generate(1000000)
        dtext, doff = synth.EnglishCorpus(dirty=True).generate(20000, first=5000000)
        with open(syn, "wb") as f:
            for t, o in ((text, off), (dtext, doff)):
                b = t.tobytes()
                for i in range(len(o) - 1):
                    f.write(b[o[i]:o[i + 1]] + b"\n")
        jobs = {
            "botchan_uni8k": "--input=%s --vocab_size=8000 --model_type=unigram --hard_vocab_limit=false" % bot,
            "botchan_bpe4k": "--input=%s --vocab_size=4000 --model_type=bpe" % bot,
            "botchan_uni2k_ud": ("--input=%s --vocab_size=2000 --model_type=unigram "
                                 "--user_defined_symbols=<sep>,Mr.,ABC --control_symbols=<ctl>") % bot,
            "botchan_bpe2k_ud": ("--input=%s --vocab_size=2000 --model_type=bpe "
                                 "--user_defined_symbols=<sep>,Mr.,ABC --control_symbols=<ctl>") % bot,
            "wagahai_uni": ("--input=%s --vocab_size=32000 --model_type=unigram --character_coverage=0.9995 "
                            "--hard_vocab_limit=false --max_sentence_length=40000") % wag,
            "syn_uni32k": ("--input=%s --vocab_size=32000 --model_type=unigram --input_sentence_size=500000 "
                           "--num_threads=8") % syn,
            "syn_bpe32k": ("--input=%s --vocab_size=32000 --model_type=bpe --input_sentence_size=200000 "
                           "--num_threads=8") % syn,
        }
        for name, args in jobs.items():
            if only and name not in only:
                continue
            prefix = os.path.join(td, name)
            ref.train(args + " --model_prefix=%s" % prefix)
            os.replace(prefix + ".model", os.path.join(OUT, name + ".model"))
            print("trained", name, os.path.getsize(os.path.join(OUT, name + ".model")), flush=True)


if __name__ == "__main__":
    main(set(sys.argv[1:]) or None)
