"""Synthetic corpora (workload generator; not on the product path).

English-like corpus of SURVEY.md 8(d) config 2/3/5 and the wagahai-style CJK
corpus of config 4, produced by tools/synth_corpus.c (built in-tree by
`__graft_entry__.build()`).  Everything is a pure function of (seed, index).
"""
import ctypes
import gzip
import os
import re
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB = os.path.join(HERE, "libsynth.so")
DATA = os.path.join(ROOT, "tests", "golden", "data")


def build(force=False):
    src = os.path.join(HERE, "synth_corpus.c")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-fopenmp", "-shared", "-fPIC", src, "-o", LIB, "-lm"])
    return LIB


class _Cfg(ctypes.Structure):
    _fields_ = [("words", ctypes.c_char_p), ("woff", ctypes.c_void_p), ("cdf", ctypes.c_void_p),
                ("nwords", ctypes.c_int32), ("lam", ctypes.c_double), ("dirty", ctypes.c_int32)]


_lib = None


def _load():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(LIB)
        _lib.synth_lengths.restype = ctypes.c_int64
        _lib.synth_resample_lengths.restype = ctypes.c_int64
    return _lib


def read_gz_lines(name):
    """Lines of a fixture text split on '\\n' (the trailing '\\r' of CRLF files
    stays in the line, as std::getline leaves it for spm_encode)."""
    with gzip.open(os.path.join(DATA, name), "rb") as f:
        raw = f.read()
    lines = raw.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    return lines


_ONSET = ["", "b", "c", "d", "f", "g", "h", "j", "k", "l", "m", "n", "p", "r", "s", "t", "v", "w",
          "z", "br", "ch", "cl", "cr", "dr", "fl", "fr", "gl", "gr", "pl", "pr", "sh", "sl", "sp",
          "st", "str", "th", "tr", "wh"]
_NUCLEUS = ["a", "e", "i", "o", "u", "ai", "au", "ea", "ee", "ie", "oo", "ou", "y"]
_CODA = ["", "", "b", "d", "g", "k", "l", "m", "n", "p", "r", "s", "t", "ck", "ct", "ld", "ll", "mp",
         "nd", "ng", "nt", "rd", "rk", "rn", "rt", "sh", "ss", "st", "th"]


def word_list(n_types=200000, seed=1234):
    """botchan's alphabetic word types (lower-cased, by descending frequency)
    followed by pseudo-words of 1-3 onset/nucleus/coda syllables."""
    text = b"\n".join(read_gz_lines("corpus_en_botchan.txt.gz")).decode("utf-8", "replace")
    freq = {}
    for w in re.findall(r"[A-Za-z]+", text):
        w = w.lower()
        freq[w] = freq.get(w, 0) + 1
    words = sorted(freq, key=lambda w: (-freq[w], w))
    seen = set(words)
    rng = np.random.RandomState(seed)
    while len(words) < n_types:
        m = 2 * (n_types - len(words)) + 1024
        k = rng.randint(1, 4, size=m)
        o = rng.randint(len(_ONSET), size=(m, 3))
        u = rng.randint(len(_NUCLEUS), size=(m, 3))
        c = rng.randint(len(_CODA), size=(m, 3))
        for i in range(m):
            w = "".join(_ONSET[o[i, j]] + _NUCLEUS[u[i, j]] + _CODA[c[i, j]] for j in range(k[i]))
            if w not in seen:
                seen.add(w)
                words.append(w)
    return words[:n_types]


class EnglishCorpus:
    def __init__(self, n_types=200000, zipf_s=1.05, lam=16.0, dirty=False, seed=1234):
        self.lib = _load()
        words = word_list(n_types, seed)
        enc = [w.encode() for w in words]
        self._bytes = b"".join(enc)
        self._woff = np.zeros(len(enc) + 1, dtype=np.int32)
        np.cumsum([len(e) for e in enc], out=self._woff[1:])
        p = 1.0 / np.arange(1, len(enc) + 1, dtype=np.float64) ** zipf_s
        self._cdf = np.cumsum(p / p.sum())
        self._cdf[-1] = 1.0
        self.cfg = _Cfg(self._bytes, self._woff.ctypes.data, self._cdf.ctypes.data, len(enc), lam,
                        1 if dirty else 0)
        self.seed = seed

    def generate(self, n, first=0):
        """(text uint8[total], offsets int64[n+1]) for sentences [first, first+n)."""
        off = np.empty(n + 1, dtype=np.int64)
        total = self.lib.synth_lengths(ctypes.byref(self.cfg), ctypes.c_uint64(self.seed),
                                       ctypes.c_int64(first), ctypes.c_int64(n),
                                       ctypes.c_void_p(off.ctypes.data))
        text = np.empty(total, dtype=np.uint8)
        self.lib.synth_fill(ctypes.byref(self.cfg), ctypes.c_uint64(self.seed), ctypes.c_int64(first),
                            ctypes.c_int64(n), ctypes.c_void_p(off.ctypes.data),
                            ctypes.c_void_p(text.ctypes.data))
        return text, off


    def word_counts(self, n, first=0):
        """The (word, freq) list of sentences [first, first + n) of the CLEAN corpus, without
        generating it: (list of word bytes, int64 freq), by descending frequency.  Exactly the
        whitespace-split tokens of generate(n, first) with their counts."""
        assert not self.cfg.dirty
        nw = self.cfg.nwords
        hist = np.zeros(3 * nw, dtype=np.int64)
        self.lib.synth_word_hist(ctypes.byref(self.cfg), ctypes.c_uint64(self.seed), ctypes.c_int64(first),
                                 ctypes.c_int64(n), ctypes.c_void_p(hist.ctypes.data))
        idx = np.flatnonzero(hist)
        order = idx[np.argsort(-hist[idx], kind="stable")]
        words = []
        for k in order:
            wi, v = divmod(int(k), 3)
            w = self._bytes[self._woff[wi]:self._woff[wi + 1]]
            if v == 1:
                w = w[:1].upper() + w[1:]
            elif v == 2:
                w = w + b"."
            words.append(w)
        return words, hist[order]


class ResampledCorpus:
    """Config 4: lines of a base text re-sampled with replacement."""

    def __init__(self, fixture="corpus_ja_wagahai.txt.gz", seed=4321):
        self.lib = _load()
        lines = read_gz_lines(fixture)
        self._bytes = b"".join(lines)
        self._loff = np.zeros(len(lines) + 1, dtype=np.int64)
        np.cumsum([len(l) for l in lines], out=self._loff[1:])
        self.nlines = len(lines)
        self.seed = seed

    def generate(self, n, first=0):
        off = np.empty(n + 1, dtype=np.int64)
        total = self.lib.synth_resample_lengths(
            ctypes.c_void_p(self._loff.ctypes.data), ctypes.c_int64(self.nlines),
            ctypes.c_uint64(self.seed), ctypes.c_int64(first), ctypes.c_int64(n),
            ctypes.c_void_p(off.ctypes.data))
        text = np.empty(total, dtype=np.uint8)
        self.lib.synth_resample_fill(
            ctypes.c_char_p(self._bytes), ctypes.c_void_p(self._loff.ctypes.data),
            ctypes.c_int64(self.nlines), ctypes.c_uint64(self.seed), ctypes.c_int64(first),
            ctypes.c_int64(n), ctypes.c_void_p(off.ctypes.data), ctypes.c_void_p(text.ctypes.data))
        return text, off


def pack_lines(lines):
    """list[bytes] -> (text uint8[], offsets int64[n+1])"""
    off = np.zeros(len(lines) + 1, dtype=np.int64)
    np.cumsum([len(l) for l in lines], out=off[1:])
    text = np.frombuffer(b"".join(lines), dtype=np.uint8).copy()
    return text, off
