#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY: build the unmodified reference into oracle/_ref/.

Compiles the reference's own encode + train sources *where they lie* under
/root/reference (g++ on the files directly, no cmake, no copy of the sources)
together with oracle/ref_shim.cc into oracle/_ref/libspm_ref.so.

Source lists follow /root/reference/src/CMakeLists.txt:63-119 (SPM_SRCS,
SPM_TRAIN_SRCS) and :18-41 (vendored protobuf-lite).  Two generated inputs:
  * gen/config.h           (config.h.in with VERSION substituted)
  * gen/normalization_rule.h -- the reference mount lacks this 4 MB data header
    (.MISSING_LARGE_BLOBS).  It only carries the precompiled charsmap blobs the
    *trainer* stamps into new models; we regenerate it from the charsmaps the
    pip `sentencepiece` wheel embeds, in the layout compile_charsmap_main.cc:94-147
    emits.  The encode path never reads it (charsmaps travel in the .model).
This only runs where /root/reference exists (the build container).  The GPU box
uses the prebuilt oracle/_ref/libspm_ref.so that travels with the snapshot.
"""
import os
import subprocess
import sys
import tempfile
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("SPM_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
GEN = os.path.join(OUT, "gen")
OBJ = os.path.join(OUT, "obj")

PB_LITE = """arena arenastring bytestream coded_stream common extension_set
generated_message_table_driven_lite generated_message_util implicit_weak_message
int128 io_win32 message_lite repeated_field status statusor stringpiece
stringprintf structurally_valid strutil time wire_format_lite zero_copy_stream
zero_copy_stream_impl_lite""".split()
SPM = """builtin_pb/sentencepiece.pb builtin_pb/sentencepiece_model.pb bpe_model
char_model error filesystem flags model_factory model_interface normalizer
sentencepiece_processor unigram_model util word_model""".split()
TRAIN = """builder unicode_script trainer_factory trainer_interface
unigram_model_trainer word_model_trainer char_model_trainer bpe_model_trainer
sentencepiece_trainer""".split()


def charsmap_blobs():
    """{rule name: precompiled charsmap bytes} from the pip sentencepiece wheel."""
    import sentencepiece as spm
    from sentencepiece import sentencepiece_model_pb2 as pb2

    blobs = {}
    corpus = os.path.join(REF, "data", "botchan.txt")
    with tempfile.TemporaryDirectory() as td:
        for name in ("nfkc", "nmt_nfkc", "nfkc_cf", "nmt_nfkc_cf"):
            prefix = os.path.join(td, name)
            spm.SentencePieceTrainer.train(
                input=corpus, model_prefix=prefix, vocab_size=300,
                normalization_rule_name=name, minloglevel=2)
            m = pb2.ModelProto()
            with open(prefix + ".model", "rb") as f:
                m.ParseFromString(f.read())
            blobs[name] = m.normalizer_spec.precompiled_charsmap
    return blobs


def write_rule_header(path):
    blobs = charsmap_blobs()
    with open(path, "w") as f:
        f.write("#ifndef NORMALIZATION_RULE_H_\n#define NORMALIZATION_RULE_H_\n#include <cstdio>\n")
        f.write("namespace sentencepiece {\nnamespace {\n")
        f.write("struct BinaryBlob {\n const char *name;\n size_t size;\n const char *data;\n};\n")
        f.write("constexpr BinaryBlob kNormalizationRules_blob[] = {\n")
        for name, data in blobs.items():
            f.write('{ "%s", %d, ' % (name, len(data)))
            for i in range(0, len(data), 20):
                f.write('"' + "".join("\\x%02X" % b for b in data[i:i + 20]) + '"\n')
            f.write("},\n")
        f.write("};\nconstexpr size_t kNormalizationRules_size = %d;\n" % len(blobs))
        f.write("}  // namespace\n}  // namespace sentencepiece\n#endif\n")


def build(with_trainer=True, force=False):
    lib = os.path.join(OUT, "libspm_ref.so")
    shim = os.path.join(HERE, "ref_shim.cc")
    if os.path.exists(lib) and not force and os.path.getmtime(lib) >= os.path.getmtime(shim):
        return lib
    if not os.path.isdir(os.path.join(REF, "src")):
        raise RuntimeError("reference sources not present at %s and no prebuilt %s" % (REF, lib))
    os.makedirs(GEN, exist_ok=True)
    os.makedirs(OBJ, exist_ok=True)
    version = open(os.path.join(REF, "VERSION")).read().strip()
    with open(os.path.join(GEN, "config.h"), "w") as f:
        f.write('#ifndef CONFIG_H_\n#define CONFIG_H_\n#define VERSION "%s"\n'
                '#define PACKAGE "sentencepiece"\n#define PACKAGE_STRING "sentencepiece"\n#endif\n' % version)
    rule_h = os.path.join(GEN, "normalization_rule.h")
    if with_trainer and not os.path.exists(rule_h):
        write_rule_header(rule_h)

    inc = ["-I" + GEN, "-I" + REF, "-I" + os.path.join(REF, "src"),
           "-I" + os.path.join(REF, "src", "builtin_pb"),
           "-I" + os.path.join(REF, "third_party", "protobuf-lite")]
    flags = ["-std=c++11", "-O3", "-fPIC", "-pthread", "-DHAVE_PTHREAD=1",
             "-Wno-sign-compare", "-Wno-deprecated-declarations", "-w"]
    srcs = [os.path.join(REF, "third_party", "protobuf-lite", s + ".cc") for s in PB_LITE]
    srcs += [os.path.join(REF, "src", s + ".cc") for s in SPM]
    srcs += [os.path.join(REF, "third_party", "absl", "strings", "string_view.cc")]
    if with_trainer:
        srcs += [os.path.join(REF, "src", s + ".cc") for s in TRAIN]
        flags.append("-DREF_WITH_TRAINER")
    srcs.append(shim)

    def compile_one(src):
        obj = os.path.join(OBJ, os.path.relpath(src, "/").replace("/", "_") + ".o")
        if (not os.path.exists(obj)) or os.path.getmtime(obj) < os.path.getmtime(src):
            subprocess.check_call(["g++"] + flags + inc + ["-c", src, "-o", obj])
        return obj

    with ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        objs = list(ex.map(compile_one, srcs))
    subprocess.check_call(["g++", "-shared", "-o", lib] + objs + ["-pthread"])
    return lib


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
