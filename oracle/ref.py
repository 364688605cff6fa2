"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/_ref/libspm_ref.so, the
unmodified reference compiled by oracle/build_ref.py.  Used as the parity
checker (tests/, smoke()) and as bench.py's reference arm / cpu_baseline.
Never imported by the product package."""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_ref", "libspm_ref.so")

_lib = None


def available():
    return os.path.exists(LIB)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            from . import build_ref
            build_ref.build()
        L = ctypes.CDLL(LIB)
        L.ref_load.restype = ctypes.c_void_p
        L.ref_load.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
        L.ref_free.argtypes = [ctypes.c_void_p]
        L.ref_last_error.restype = ctypes.c_char_p
        L.ref_set_encode_extra_options.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
        L.ref_set_vocabulary.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int]
        L.ref_piece_size.argtypes = [ctypes.c_void_p]
        L.ref_piece_type.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.ref_encode_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                       ctypes.c_int64, ctypes.c_int, ctypes.c_void_p,
                                       ctypes.c_int64, ctypes.c_void_p]
        L.ref_encode_count.restype = ctypes.c_int64
        L.ref_encode_count.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                       ctypes.c_int64, ctypes.c_int]
        L.ref_normalize.restype = ctypes.c_int64
        L.ref_encode_proto_batch.restype = ctypes.c_int64
        L.ref_encode_proto_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                             ctypes.c_int, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p,
                                             ctypes.c_int]
        L.ref_decode_batch.restype = ctypes.c_int64
        L.ref_decode_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                       ctypes.c_int, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p,
                                       ctypes.c_int]
        L.ref_normalize.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int64,
                                    ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
        if hasattr(L, "ref_train"):
            L.ref_train.argtypes = [ctypes.c_char_p]
        if hasattr(L, "ref_estep"):
            L.ref_estep.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p,
                                    ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p,
                                    ctypes.c_void_p, ctypes.c_void_p]
        if hasattr(L, "ref_estep64"):
            L.ref_estep64.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p,
                                      ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p,
                                      ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        _lib = L
    return _lib


class RefProcessor:
    """The reference SentencePieceProcessor (Load / Encode(ids) batch loop)."""

    def __init__(self, model_bytes):
        self.L = lib()
        self.h = self.L.ref_load(model_bytes, len(model_bytes))
        if not self.h:
            raise RuntimeError("reference Load failed: %s" % self.L.ref_last_error().decode())

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_free(self.h)
            self.h = None

    @classmethod
    def from_file(cls, path):
        with open(path, "rb") as f:
            return cls(f.read())

    def set_encode_extra_options(self, opts):
        rc = self.L.ref_set_encode_extra_options(self.h, opts.encode())
        if rc:
            raise ValueError(self.L.ref_last_error().decode())

    def set_vocabulary(self, pieces):
        if pieces is None:
            rc = self.L.ref_set_vocabulary(self.h, None, -1)
        else:
            buf = b"".join(p.encode() + b"\0" for p in pieces)
            rc = self.L.ref_set_vocabulary(self.h, buf, len(pieces))
        if rc:
            raise ValueError(self.L.ref_last_error().decode())

    def piece_size(self):
        return self.L.ref_piece_size(self.h)

    def encode_batch(self, text, off, threads=1):
        """text uint8[], off int64[n+1] -> (ids int32[], id_off int64[n+1])"""
        text = np.ascontiguousarray(text, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.int64)
        n = len(off) - 1
        id_off = np.zeros(n + 1, dtype=np.int64)
        # ids capacity bound: one id per byte + bos/eos
        cap = int(off[-1] - off[0]) + 2 * n + 16
        ids = np.empty(cap, dtype=np.int32)
        rc = self.L.ref_encode_batch(self.h, text.ctypes.data, off.ctypes.data, n, threads,
                                     ids.ctypes.data, cap, id_off.ctypes.data)
        if rc:
            raise RuntimeError("reference Encode status %d" % rc)
        return ids[: id_off[n]].copy(), id_off

    def encode_count(self, text, off, threads=1):
        text = np.ascontiguousarray(text, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.int64)
        return self.L.ref_encode_count(self.h, text.ctypes.data, off.ctypes.data, len(off) - 1,
                                       threads)

    def encode(self, s):
        b = s.encode() if isinstance(s, str) else bytes(s)
        text = np.frombuffer(b, dtype=np.uint8)
        ids, _ = self.encode_batch(text, np.array([0, len(b)], dtype=np.int64))
        return ids.tolist()

    def encode_proto_batch(self, text, off, threads=1):
        """EncodeAsSerializedProto per sentence -> (proto bytes uint8[], proto_off int64[n+1])."""
        text = np.ascontiguousarray(text, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.int64)
        n = len(off) - 1
        poff = np.zeros(n + 1, dtype=np.int64)
        tot = self.L.ref_encode_proto_batch(self.h, text.ctypes.data, off.ctypes.data, n, threads, None, 0,
                                            poff.ctypes.data, 0)
        out = np.empty(max(int(tot), 1), dtype=np.uint8)
        self.L.ref_encode_proto_batch(self.h, text.ctypes.data, off.ctypes.data, n, threads, out.ctypes.data,
                                      int(tot), poff.ctypes.data, 0)
        return out[: int(tot)], poff

    def encode_pieces_count(self, text, off, threads=1):
        text = np.ascontiguousarray(text, dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.int64)
        return self.L.ref_encode_proto_batch(self.h, text.ctypes.data, off.ctypes.data, len(off) - 1, threads,
                                             None, 0, None, 1)

    def decode_batch(self, ids, id_off, threads=1):
        """Decode(ids) per sentence -> (text uint8[], text_off int64[n+1])."""
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        id_off = np.ascontiguousarray(id_off, dtype=np.int64)
        n = len(id_off) - 1
        toff = np.zeros(n + 1, dtype=np.int64)
        tot = self.L.ref_decode_batch(self.h, ids.ctypes.data, id_off.ctypes.data, n, threads, None, 0,
                                      toff.ctypes.data, 0)
        out = np.empty(max(int(tot), 1), dtype=np.uint8)
        self.L.ref_decode_batch(self.h, ids.ctypes.data, id_off.ctypes.data, n, threads, out.ctypes.data, int(tot),
                                toff.ctypes.data, 0)
        return out[: int(tot)], toff

    def decode_count(self, ids, id_off, threads=1):
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        id_off = np.ascontiguousarray(id_off, dtype=np.int64)
        return self.L.ref_decode_batch(self.h, ids.ctypes.data, id_off.ctypes.data, len(id_off) - 1, threads,
                                       None, 0, None, 1)

    def normalize(self, s):
        b = s.encode() if isinstance(s, str) else bytes(s)
        cap = 40 * len(b) + 64
        out = ctypes.create_string_buffer(cap)
        n2o = np.zeros(cap + 1, dtype=np.int64)
        n = self.L.ref_normalize(self.h, b, len(b), out, cap, n2o.ctypes.data)
        if n < 0:
            raise RuntimeError("ref_normalize failed")
        return out.raw[:n], n2o[: n + 1].copy() if n else n2o[:0]


def train(args):
    rc = lib().ref_train(args.encode())
    if rc:
        raise RuntimeError("reference spm_train failed: %s" % lib().ref_last_error().decode())


def estep(pieces, scores, words, freq):
    """Reference E-step (TrainerModel + Lattice::PopulateMarginal + Viterbi count, the loop of
    Trainer::RunEStep, src/unigram_model_trainer.cc:256-267) over (word, freq) pairs.
    pieces / words: lists of bytes without NUL.  Returns (expected f32[V], logz f32[n], obj, ntokens)."""
    L = lib()
    if not hasattr(L, "ref_estep"):
        raise RuntimeError("reference library was built without the trainer")
    scores = np.ascontiguousarray(scores, dtype=np.float32)
    freq = np.ascontiguousarray(freq, dtype=np.int64)
    pb = b"".join(p + b"\0" for p in pieces)
    wb = b"".join(w + b"\0" for w in words)
    expected = np.zeros(len(pieces), dtype=np.float32)
    logz = np.zeros(len(words), dtype=np.float32)
    obj = ctypes.c_double(0)
    nt = ctypes.c_int64(0)
    rc = L.ref_estep(pb, scores.ctypes.data, len(pieces), wb, freq.ctypes.data, len(words),
                     expected.ctypes.data, logz.ctypes.data, ctypes.byref(obj), ctypes.byref(nt))
    if rc:
        raise RuntimeError("ref_estep failed")
    return expected, logz, obj.value, nt.value


def estep64(pieces, scores, words, freq):
    """estep() plus expected64 = the float64 re-accumulation of the reference's own per-word
    posteriors: (expected f32[V], logz f32[n], obj, ntokens, expected64 f64[V])."""
    L = lib()
    if not hasattr(L, "ref_estep64"):
        raise RuntimeError("reference library was built without ref_estep64")
    scores = np.ascontiguousarray(scores, dtype=np.float32)
    freq = np.ascontiguousarray(freq, dtype=np.int64)
    pb = b"".join(p + b"\0" for p in pieces)
    wb = b"".join(w + b"\0" for w in words)
    expected = np.zeros(len(pieces), dtype=np.float32)
    expected64 = np.zeros(len(pieces), dtype=np.float64)
    logz = np.zeros(len(words), dtype=np.float32)
    obj = ctypes.c_double(0)
    nt = ctypes.c_int64(0)
    rc = L.ref_estep64(pb, scores.ctypes.data, len(pieces), wb, freq.ctypes.data, len(words),
                       expected.ctypes.data, logz.ctypes.data, ctypes.byref(obj), ctypes.byref(nt),
                       expected64.ctypes.data)
    if rc:
        raise RuntimeError("ref_estep64 failed")
    return expected, logz, obj.value, nt.value, expected64
