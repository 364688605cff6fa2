"""TEST INFRASTRUCTURE ONLY: ctypes binding of the CPU restatement
(oracle/spm_oracle.cc -> oracle/_build/libspm_oracle.so).  Checker for tests/,
smoke() and bench.py's cpu_baseline; never imported by the product package."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "spm_oracle.cc")
LIB = os.path.join(HERE, "_build", "libspm_oracle.so")

OPT_CODES = {"bos": 1, "eos": 2, "reverse": 3}


def pack_options(opts):
    """'bos:eos:reverse' -> 2-bit codes packed low bits first (see include/spm_cuda.h)."""
    v = 0
    if not opts:
        return 0
    for k, name in enumerate(opts.split(":")):
        if name not in OPT_CODES:
            raise ValueError('option "%s" is not available.' % name)
        if k >= 16:
            raise ValueError("too many extra options")
        v |= OPT_CODES[name] << (2 * k)
    return v


def build(force=False):
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++14", "-shared", "-fPIC", SRC, "-o", LIB])
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(LIB)
        L.oracle_load.restype = ctypes.c_void_p
        L.oracle_load.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
        L.oracle_free.argtypes = [ctypes.c_void_p]
        L.oracle_piece_size.argtypes = [ctypes.c_void_p]
        L.oracle_model_type.argtypes = [ctypes.c_void_p]
        L.oracle_set_piece_types.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        L.oracle_get_piece_types.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.oracle_normalize.restype = ctypes.c_int64
        L.oracle_normalize.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int64,
                                       ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
        L.oracle_encode_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                          ctypes.c_int64, ctypes.c_uint32, ctypes.c_void_p,
                                          ctypes.c_int64, ctypes.c_void_p]
        L.oracle_encode_pieces_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                                 ctypes.c_uint32] + [ctypes.c_void_p] * 5 + [
                                                 ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                                 ctypes.c_void_p]
        L.oracle_decode_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                          ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
        _lib = L
    return _lib


class OracleProcessor:
    def __init__(self, model_bytes):
        self.L = lib()
        self.h = self.L.oracle_load(model_bytes, len(model_bytes))
        if not self.h:
            raise RuntimeError("oracle_load failed")
        self.opts = 0

    def __del__(self):
        if getattr(self, "h", None):
            self.L.oracle_free(self.h)
            self.h = None

    @classmethod
    def from_file(cls, path):
        with open(path, "rb") as f:
            return cls(f.read())

    def set_encode_extra_options(self, opts):
        self.opts = pack_options(opts)

    def piece_size(self):
        return self.L.oracle_piece_size(self.h)

    def piece_types(self):
        t = np.zeros(self.piece_size(), dtype=np.uint8)
        self.L.oracle_get_piece_types(self.h, t.ctypes.data)
        return t

    def set_piece_types(self, types):
        t = np.ascontiguousarray(types, dtype=np.uint8)
        self.L.oracle_set_piece_types(self.h, t.ctypes.data, len(t))

    def encode_batch(self, text, off):
        text = np.ascontiguousarray(text, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.int64)
        n = len(off) - 1
        id_off = np.zeros(n + 1, dtype=np.int64)
        cap = int(off[-1] - off[0]) + 34 * n + 64
        ids = np.empty(cap, dtype=np.int32)
        rc = self.L.oracle_encode_batch(self.h, text.ctypes.data, off.ctypes.data, n, self.opts,
                                        ids.ctypes.data, cap, id_off.ctypes.data)
        if rc == 8:
            raise RuntimeError("oracle ids capacity exceeded")
        self.last_status = rc
        return ids[: id_off[n]].copy(), id_off

    def encode(self, s):
        b = s.encode() if isinstance(s, str) else bytes(s)
        ids, _ = self.encode_batch(np.frombuffer(b, dtype=np.uint8), np.array([0, len(b)], dtype=np.int64))
        return ids.tolist()

    def normalize(self, s):
        b = s.encode() if isinstance(s, str) else bytes(s)
        cap = 40 * len(b) + 64
        out = ctypes.create_string_buffer(cap)
        n2o = np.zeros(cap + 1, dtype=np.int64)
        n = self.L.oracle_normalize(self.h, b, len(b), out, cap, n2o.ctypes.data)
        if n < 0:
            raise RuntimeError("oracle_normalize capacity")
        return out.raw[:n], (n2o[: n + 1].copy() if n else n2o[:0])

    def encode_pieces_batch(self, text, off):
        """SentencePieceText view as CSR arrays (layout of spm_cuda_encode_batch_pieces):
        dict(ids, begin, end, nbegin, nend, id_off, norm, norm_off)."""
        text = np.ascontiguousarray(text, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.int64)
        n = len(off) - 1
        cap = int(off[-1] - off[0]) + 34 * n + 64
        ncap = 40 * int(off[-1] - off[0]) + 8 * n + 64
        a = {k: np.empty(cap, dtype=np.int32) for k in ("ids", "begin", "end", "nbegin", "nend")}
        id_off = np.zeros(n + 1, dtype=np.int64)
        norm = np.empty(ncap, dtype=np.uint8)
        norm_off = np.zeros(n + 1, dtype=np.int64)
        rc = self.L.oracle_encode_pieces_batch(self.h, text.ctypes.data, off.ctypes.data, n, self.opts,
                                               a["ids"].ctypes.data, a["begin"].ctypes.data, a["end"].ctypes.data,
                                               a["nbegin"].ctypes.data, a["nend"].ctypes.data, cap,
                                               id_off.ctypes.data, norm.ctypes.data, ncap, norm_off.ctypes.data)
        if rc == 8:
            raise RuntimeError("oracle capacity exceeded")
        self.last_status = rc
        out = {k: v[: id_off[n]].copy() for k, v in a.items()}
        out.update(id_off=id_off, norm=norm[: norm_off[n]].copy(), norm_off=norm_off)
        return out

    def decode_batch(self, ids, id_off):
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        id_off = np.ascontiguousarray(id_off, dtype=np.int64)
        n = len(id_off) - 1
        cap = 64 * len(ids) + 64
        text = np.empty(cap, dtype=np.uint8)
        toff = np.zeros(n + 1, dtype=np.int64)
        rc = self.L.oracle_decode_batch(self.h, ids.ctypes.data, id_off.ctypes.data, n, text.ctypes.data, cap,
                                        toff.ctypes.data)
        if rc:
            raise RuntimeError("oracle_decode_batch: code %d" % rc)
        return text[: toff[n]].copy(), toff
