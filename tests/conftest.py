import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

MODELS = os.path.join(ROOT, "tests", "golden", "models")
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def model_bytes(name):
    with open(os.path.join(MODELS, name + ".model"), "rb") as f:
        return f.read()


@pytest.fixture(scope="session")
def spm_b200():
    import spm_b200_pkg
    return spm_b200_pkg.load()


@pytest.fixture(scope="session")
def corpora():
    """Small seeded corpora shared by the parity tests: name -> (text u8, offsets i64)."""
    from tools import synth
    bot = synth.pack_lines(synth.read_gz_lines("corpus_en_botchan.txt.gz"))
    wag = synth.pack_lines(synth.read_gz_lines("corpus_ja_wagahai.txt.gz"))
    en = synth.EnglishCorpus().generate(20000)
    dirty = synth.EnglishCorpus(dirty=True).generate(20000, first=777)
    edge = synth.pack_lines(EDGE_SENTENCES)
    return {"botchan": bot, "wagahai": wag, "syn_en": en, "syn_dirty": dirty, "edge": edge}


# empty / whitespace-only / malformed / expansion-heavy / user-symbol inputs
EDGE_SENTENCES = [
    b"", b" ", b"   ", b"\t", b"\xe3\x80\x80", b"a", b" a ", b"  a  b  ", b"\xe2\x96\x81", b"\xe2\x96\x81\xe2\x96\x81a\xe2\x96\x81",
    b"abc\x80xy", b"ab\xe3\x81xy", b"\xff\xfe\xfd", b"\xf0\x9f\x98", b"\xc0\xaf", b"\xed\xa0\x80", b"\xef\xbf\xbd",
    b"\x00", b"a\x00b", b"\x01\x02\x7f", "ｸﾞｰｸﾞﾙ".encode(), " ｸﾞｰｸﾞﾙ ".encode(), "㍿".encode(), "ﷺ".encode() * 5,
    "ﷺ ﷺ".encode(), "①②③".encode(), "½ ¼".encode(), "Ａ̀".encode(), "Å".encode(), "éé".encode(),
    b"www", b"wwww", b"wwwww www", b"aaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaa",
    b"Mr. Smith <sep> ABC<sep>Mr.", b"<sep>", b"<ctl>", b"<s>hello</s>", b"<unk>", b"hello<unk><unk>world",
    "こんにちは世界".encode(), "吾輩は猫である。名前はまだ無い。".encode(), "\U0001F600\U0001F600 emoji \U0001F4A9".encode(),
    "Ünïcödé tëxt wïth äccénts".encode(), b"tab\tseparated\t\tvalues", b"trailing space   ", b"   leading space",
    b"x" * 5000, ("あ" * 3000).encode(), b"The quick brown fox jumps over the lazy dog.",
    b"1234567890 !@#$%^&*() -=_+ []{} ;':\",./<>?", "«quoted» — dash … ellipsis".encode(),
]


def near_tie_model(kmax=20, eps=3e-7):
    """A unigram model whose pieces a, aa, ... a^kmax have near-additive scores: in a long run of
    'a' more than 8 (the exact tier's first capacity) DISTINCT near-tied candidates end at one
    position, so the sentence must take the retry with room for every candidate."""
    import protowire as pw
    ps = pw.base_pieces() + [("\u2581", -2.0)]
    ps += [("a" * k, -float(k) - eps * k * k) for k in range(1, kmax + 1)]
    ps += [("b", -3.0), ("\u2581b", -3.5)]
    return pw.model(ps)


NEAR_TIE_SENTENCES = [b"a" * n for n in (1, 5, 13, 40, 100, 300, 1000, 5000)] + \
    [b"aaaa b " + b"a" * 77 + b" ab" * 5, b"b" + b"a" * 30 + b"b", b"", b"b b b"]


def _build_lib(src_files, out, extra):
    os.makedirs(os.path.dirname(out), exist_ok=True)
    newest = max(os.path.getmtime(s) for s in src_files)
    if not os.path.exists(out) or os.path.getmtime(out) < newest:
        subprocess.check_call(["g++", "-O2", "-std=c++14", "-shared", "-fPIC"] + extra + ["-o", out])
    return out


@pytest.fixture(scope="session")
def hostcore():
    """tests/host_core.cc: the kernels' __host__ __device__ core compiled for the CPU (test-only)."""
    csrc = os.path.join(ROOT, "sentencepiece-comments_b200", "csrc")
    srcs = [os.path.join(ROOT, "tests", "host_core.cc"), os.path.join(csrc, "spm_model.cc")]
    deps = srcs + [os.path.join(csrc, "spm_core.cuh"), os.path.join(csrc, "spm_fast.cuh"), os.path.join(csrc, "spm_memo.cuh"), os.path.join(csrc, "spm_estep.cuh"),
                   os.path.join(csrc, "spm_model.h")]
    out = os.path.join(ROOT, "tests", "_build", "libspm_hostcore.so")
    _build_lib(deps, out, ["-I", csrc] + srcs)
    L = ctypes.CDLL(out)
    L.hc_load.restype = ctypes.c_void_p
    L.hc_load.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p]
    L.hc_free.argtypes = [ctypes.c_void_p]
    L.hc_expansion.argtypes = [ctypes.c_void_p]
    L.hc_fast_flags.argtypes = [ctypes.c_void_p]
    L.hc_set_piece_types.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    L.hc_normalize.restype = ctypes.c_int64
    L.hc_normalize.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64]
    L.hc_normalize_fast.restype = ctypes.c_int64
    L.hc_normalize_fast.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64,
                                    ctypes.c_void_p]
    L.hc_normalize_chunked.restype = ctypes.c_int64
    L.hc_normalize_chunked.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p,
                                       ctypes.c_int64]
    L.hc_encode_batch.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int64, ctypes.c_uint32, ctypes.c_int,
                                                         ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p,
                                                         ctypes.c_void_p]
    L.hc_lse_table_error.restype = ctypes.c_double
    L.hc_lse_table_error.argtypes = [ctypes.c_int64]
    L.hc_estep.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                           ctypes.c_void_p, ctypes.c_int64, ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p,
                           ctypes.c_void_p]
    return L


class HostCore:
    def __init__(self, L, mb):
        self.L = L
        code = ctypes.c_int(0)
        self.h = L.hc_load(mb, len(mb), ctypes.byref(code))
        self.code = code.value
        if not self.h:
            raise RuntimeError("hc_load failed with code %d" % code.value)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.hc_free(self.h)

    def fast_flags(self):
        return self.L.hc_fast_flags(self.h)

    def set_piece_types(self, t):
        t = np.ascontiguousarray(t, dtype=np.uint8)
        self.L.hc_set_piece_types(self.h, t.ctypes.data, len(t))

    def normalize(self, b):
        cap = self.L.hc_expansion(self.h) * len(b) + 8
        out = ctypes.create_string_buffer(cap)
        n = self.L.hc_normalize(self.h, b, len(b), out, cap)
        assert n >= 0
        return out.raw[:n]

    def normalize_fast(self, b):
        """fast-tier normaliser; None when the sentence is routed to the general tier."""
        cap = 3 * len(b) + 80
        out = ctypes.create_string_buffer(cap)
        nc = ctypes.c_int32(0)
        n = self.L.hc_normalize_fast(self.h, b, len(b), out, cap, ctypes.byref(nc))
        assert n >= -1, n
        return None if n < 0 else (out.raw[:n], nc.value)

    def normalize_chunked(self, b, chunk):
        """k_normalize_long's piecewise scheme; None when the line is routed to the general tier."""
        cap = 3 * len(b) + 80
        out = ctypes.create_string_buffer(cap)
        n = self.L.hc_normalize_chunked(self.h, b, len(b), chunk, out, cap)
        assert n >= -1, n
        return None if n < 0 else out.raw[:n]

    def encode_batch(self, text, off, opts=0, capc=4):
        text = np.ascontiguousarray(text, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.int64)
        n = len(off) - 1
        cap = int(off[-1]) * 4 + 40 * n + 64
        ids = np.empty(cap, dtype=np.int32)
        io = np.zeros(n + 1, dtype=np.int64)
        st = np.zeros(max(n, 1), dtype=np.uint32)
        rc = self.L.hc_encode_batch(self.h, text.ctypes.data, off.ctypes.data, n, opts, capc,
                                    ids.ctypes.data, cap, io.ctypes.data, st.ctypes.data)
        assert rc == 0, rc
        return ids[: io[n]].copy(), io, st[:n]

    def encode_pieces_batch(self, text, off, opts=0):
        text = np.ascontiguousarray(text, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.int64)
        n = len(off) - 1
        cap = int(off[-1] - off[0]) * 4 + 40 * n + 64
        ncap = self.L.hc_expansion(self.h) * int(off[-1] - off[0]) + 16 * n + 64
        a = {k: np.empty(cap, dtype=np.int32) for k in ("ids", "begin", "end", "nbegin", "nend")}
        io = np.zeros(n + 1, dtype=np.int64)
        norm = np.empty(ncap, dtype=np.uint8)
        no = np.zeros(n + 1, dtype=np.int64)
        flagged = ctypes.c_int64(0)     # sentences spans_from_ids hands back to the exact tier
        plain = ctypes.c_int64(0)       # sentences the plain-sentence path settles (checked against the exact spans)
        rc = self.L.hc_encode_pieces_batch(
            ctypes.c_void_p(self.h), ctypes.c_void_p(text.ctypes.data), ctypes.c_void_p(off.ctypes.data),
            ctypes.c_int64(n), ctypes.c_uint32(opts), *[ctypes.c_void_p(a[k].ctypes.data) for k in
                                                        ("ids", "begin", "end", "nbegin", "nend")],
            ctypes.c_int64(cap), ctypes.c_void_p(io.ctypes.data), ctypes.c_void_p(norm.ctypes.data),
            ctypes.c_int64(ncap), ctypes.c_void_p(no.ctypes.data), ctypes.byref(flagged), ctypes.byref(plain))
        assert rc == 0, rc
        out = {k: v[: io[n]].copy() for k, v in a.items()}
        out.update(id_off=io, norm=norm[: no[n]].copy(), norm_off=no, flagged=flagged.value, plain=plain.value)
        return out


@pytest.fixture(scope="session")
def make_hostcore(hostcore):
    return lambda mb: HostCore(hostcore, mb)


def checker(mb):
    """The parity checker: the unmodified reference (oracle/_ref) when its
    library is present, else the CPU restatement (oracle/spm_oracle.cc)."""
    from oracle import port, ref
    if ref.available():
        return ref.RefProcessor(mb), "reference"
    return port.OracleProcessor(mb), "port"
