"""B200-native batch encoder for SentencePiece models.

Python mirror of the reference's `sentencepiece.SentencePieceProcessor`
(python/sentencepiece.i:138-276 wraps src/sentencepiece_processor.h:176-447)
for the encode path: same method names, argument meaning and error behaviour
(the SWIG build raises on a non-OK util::Status, sentencepiece_processor.h:316-320),
plus the batch entry points the GPU path is built for.  Everything goes through
the C ABI of include/spm_cuda.h (libspm_cuda.so); there is no CPU fallback --
importing works without a GPU, loading a model does not.
"""
import ctypes
import os

import numpy as np

from . import build as _build
from . import spt

_HERE = os.path.dirname(os.path.abspath(__file__))
# SPM_B200_LIB_VARIANT: a differently compiled libspm_cuda.so (same sources, other build-time knobs) for A/B measurements
LIB_PATH = os.environ.get("SPM_B200_LIB_VARIANT") or _build.LIB

HOST, DEVICE = 0, 1
_OPT_CODES = {"bos": 1, "eos": 2, "reverse": 3}
NORMAL, UNKNOWN, CONTROL, USER_DEFINED, UNUSED = 1, 2, 3, 4, 5

_lib = None


def lib():
    """The C-ABI library; fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libspm_cuda.so is not built: run `python __graft_entry__.py` "
                               "(there is no CPU fallback)")
        L = ctypes.CDLL(LIB_PATH)
        vp, i64, i32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
        L.spm_cuda_model_load.argtypes = [ctypes.c_char_p, ctypes.c_size_t, i32, ctypes.POINTER(vp)]
        L.spm_cuda_model_free.argtypes = [vp]
        L.spm_cuda_model_set_piece_types.argtypes = [vp, vp, i32]
        L.spm_cuda_model_set_option.argtypes = [vp, ctypes.c_char_p, i64]
        L.spm_cuda_model_get_option.argtypes = [vp, ctypes.c_char_p, ctypes.POINTER(i64)]
        L.spm_cuda_model_release_workspace.argtypes = [vp]
        L.spm_cuda_encode_batch.argtypes = [vp, vp, vp, i64, ctypes.c_uint32, vp, i64, vp, i32, vp]
        L.spm_cuda_model_piece_size.argtypes = [vp]
        L.spm_cuda_model_piece.argtypes = [vp, i32, ctypes.POINTER(i32)]
        L.spm_cuda_model_piece.restype = ctypes.c_void_p
        L.spm_cuda_model_score.argtypes = [vp, i32]
        L.spm_cuda_model_score.restype = ctypes.c_float
        L.spm_cuda_model_piece_type.argtypes = [vp, i32]
        L.spm_cuda_model_piece_to_id.argtypes = [vp, ctypes.c_char_p, ctypes.c_size_t]
        for f in ("type", "unk_id", "bos_id", "eos_id", "pad_id"):
            getattr(L, "spm_cuda_model_" + f).argtypes = [vp]
        L.spm_cuda_model_check_extra_options.argtypes = [vp, ctypes.c_uint32]
        L.spm_cuda_last_stats.argtypes = [vp, vp, i32]
        L.spm_cuda_last_error.restype = ctypes.c_char_p
        L.spm_cuda_estep_create.argtypes = [ctypes.c_char_p, vp, i32, i32, ctypes.POINTER(vp)]
        L.spm_cuda_estep_free.argtypes = [vp]
        L.spm_cuda_estep_vocab_size.argtypes = [vp]
        L.spm_cuda_estep_run.argtypes = [vp, vp, vp, vp, i64, ctypes.c_double, vp, vp, vp, i32, vp]
        L.spm_cuda_split_lines.argtypes = [vp, vp, i64, vp, vp, i64, ctypes.POINTER(i64), i32, vp]
        L.spm_cuda_split_lines_noalloc.argtypes = [vp, vp, i64, vp, vp, i64, ctypes.POINTER(i64), vp]
        L.spm_cuda_encode_batch_pieces.argtypes = [vp, vp, vp, i64, ctypes.c_uint32, vp, vp, vp, vp, vp, i64, vp,
                                                   vp, i64, vp, i32, vp]
        L.spm_cuda_decode_batch.argtypes = [vp, vp, vp, i64, vp, i64, vp, i32, vp]
        L.spm_cuda_csr_max_len.argtypes = [vp, vp, i64, ctypes.POINTER(i64), vp]
        L.spm_cuda_csr_to_dense.argtypes = [vp, vp, vp, i64, i32, vp, i64, vp, vp]
        L.spm_cuda_csr_to_sparse.argtypes = [vp, vp, i64, vp, vp]
        _lib = L
    return _lib


class StatusError(RuntimeError):
    """A non-OK util::Status (code = util::error::Code)."""

    def __init__(self, code, message):
        super().__init__("%s (code %d)" % (message, code))
        self.code = code


def _check(rc):
    if rc != 0:
        raise StatusError(rc, lib().spm_cuda_last_error().decode("utf-8", "replace"))


def pack_extra_options(opts):
    """'bos:eos:reverse' -> packed option list (include/spm_cuda.h); validates
    names as ParseExtraOptions does (src/sentencepiece_processor.cc:991-1023)."""
    v = 0
    if not opts:
        return 0
    names = opts.split(":")
    if len(names) > 16:
        raise StatusError(3, "too many extra options")
    for k, name in enumerate(names):
        if name not in _OPT_CODES:
            raise StatusError(13, 'option "%s" is not available.' % name)
        v |= _OPT_CODES[name] << (2 * k)
    return v


class SentencePieceProcessor:
    def __init__(self, model_file=None, device=0):
        self._h = None
        self._device = device
        self._opts = 0
        self._base_types = None
        if model_file is not None:
            self.Load(model_file)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().spm_cuda_model_free(self._h)
            self._h = None

    # ---- loading (sentencepiece_processor.h:183-203)
    def Load(self, filename):
        try:
            with open(filename, "rb") as f:
                data = f.read()
        except OSError:
            raise StatusError(5, '"%s": No such file or directory' % filename)
        return self.LoadFromSerializedProto(data)

    def LoadFromSerializedProto(self, serialized):
        h = ctypes.c_void_p()
        _check(lib().spm_cuda_model_load(serialized, len(serialized), self._device, ctypes.byref(h)))
        if self._h:
            lib().spm_cuda_model_free(self._h)
        self._h = h
        n = self.GetPieceSize()
        self._base_types = np.array([lib().spm_cuda_model_piece_type(h, i) for i in range(n)], dtype=np.uint8)
        return True

    def _need(self):
        if not self._h:
            raise StatusError(9, "Model is not initialized.")

    # ---- options (sentencepiece_processor.h:210-234)
    def SetEncodeExtraOptions(self, extra_option):
        self._need()
        v = pack_extra_options(extra_option)
        _check(lib().spm_cuda_model_check_extra_options(self._h, v))
        self._opts = v
        return True

    def SetVocabulary(self, valid_vocab):
        """src/sentencepiece_processor.cc:203-231: pieces outside `valid_vocab`
        (and longer than one character) become UNUSED."""
        self._need()
        vocab = set(v.encode() if isinstance(v, str) else bytes(v) for v in valid_vocab)
        types = self._base_types.copy()
        for i in range(len(types)):
            if types[i] in (CONTROL, UNKNOWN, USER_DEFINED):
                continue
            p = self._piece_bytes(i)
            one = (1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 3, 4)[p[0] >> 4]
            types[i] = NORMAL if (p in vocab or one == len(p)) else UNUSED
        _check(lib().spm_cuda_model_set_piece_types(self._h, types.ctypes.data, len(types)))
        self._types = types
        return True

    def ResetVocabulary(self):
        self._need()
        types = self._base_types.copy()
        types[types == UNUSED] = NORMAL
        _check(lib().spm_cuda_model_set_piece_types(self._h, types.ctypes.data, len(types)))
        return True

    # ---- per-handle knobs (include/spm_cuda.h: spm_cuda_model_set_option; tests / measurements)
    def set_option(self, name, value):
        self._need()
        _check(lib().spm_cuda_model_set_option(self._h, name.encode(), int(value)))

    def get_option(self, name):
        self._need()
        v = ctypes.c_int64(0)
        _check(lib().spm_cuda_model_get_option(self._h, name.encode(), ctypes.byref(v)))
        return v.value

    def release_workspace(self):
        self._need()
        _check(lib().spm_cuda_model_release_workspace(self._h))

    # ---- encode (sentencepiece_processor.h:244, 333-335)
    def Encode(self, input):
        return self.EncodeAsIds(input)

    def EncodeAsIds(self, input):
        b = input.encode("utf-8") if isinstance(input, str) else bytes(input)
        ids, _ = self.encode_packed(np.frombuffer(b, dtype=np.uint8), np.array([0, len(b)], dtype=np.int64))
        return ids.tolist()

    def encode_batch(self, sentences):
        """list of str/bytes -> (ids int32[total], id_offsets int64[n+1])."""
        bs = [s.encode("utf-8") if isinstance(s, str) else bytes(s) for s in sentences]
        off = np.zeros(len(bs) + 1, dtype=np.int64)
        if bs:
            np.cumsum([len(b) for b in bs], out=off[1:])
        text = np.frombuffer(b"".join(bs), dtype=np.uint8)
        return self.encode_packed(text, off)

    def encode_packed(self, text, offsets, ids_capacity=None, stream=None):
        """Host CSR in, host CSR out.  text: uint8[], offsets: int64[n+1]."""
        self._need()
        text = np.ascontiguousarray(text, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        n = len(offsets) - 1
        id_off = np.zeros(n + 1, dtype=np.int64)
        cap = ids_capacity if ids_capacity is not None else max(64, (int(offsets[-1] - offsets[0]) // 2) + 4 * n)
        while True:
            ids = np.empty(cap, dtype=np.int32)
            rc = lib().spm_cuda_encode_batch(self._h, text.ctypes.data, offsets.ctypes.data, n, self._opts,
                                             ids.ctypes.data, cap, id_off.ctypes.data, HOST, stream)
            if rc == 8 and ids_capacity is None:
                cap = int(id_off[n])
                continue
            _check(rc)
            return ids[: id_off[n]], id_off

    def encode_device(self, text, offsets, ids, id_offsets, stream=None):
        """Device CSR in/out on torch CUDA tensors (uint8, int64, int32, int64);
        runs on torch's current stream unless `stream` (a cudaStream_t int) is given."""
        import torch
        self._need()
        if stream is None:
            stream = torch.cuda.current_stream(text.device).cuda_stream
        n = offsets.numel() - 1
        rc = lib().spm_cuda_encode_batch(self._h, text.data_ptr(), offsets.data_ptr(), n, self._opts,
                                         ids.data_ptr(), ids.numel(), id_offsets.data_ptr(), DEVICE, stream)
        _check(rc)

    def encode_pinned(self, text, offsets, ids, id_offsets, stream=None):
        """Host CSR in/out on caller-owned (ideally pinned) torch/numpy buffers; no allocation."""
        self._need()
        ptr = lambda a: a.data_ptr() if hasattr(a, "data_ptr") else a.ctypes.data
        n = (offsets.numel() if hasattr(offsets, "numel") else len(offsets)) - 1
        cap = ids.numel() if hasattr(ids, "numel") else len(ids)
        rc = lib().spm_cuda_encode_batch(self._h, ptr(text), ptr(offsets), n, self._opts, ptr(ids), cap,
                                         ptr(id_offsets), HOST, stream)
        _check(rc)

    def split_lines(self, buf, stream=None):
        """A raw file buffer (bytes / uint8 array) -> (text uint8[], offsets int64[n+1]): the
        lines std::getline hands to spm_encode (src/spm_encode_main.cc:183-193), split on the GPU."""
        import numpy as np
        self._need()
        buf = np.frombuffer(buf, dtype=np.uint8) if isinstance(buf, (bytes, bytearray, memoryview)) else \
            np.ascontiguousarray(buf, dtype=np.uint8)
        n = ctypes.c_int64(0)
        cap = max(16, len(buf) // 16)
        while True:
            text = np.empty(max(len(buf), 1), dtype=np.uint8)
            off = np.empty(cap + 1, dtype=np.int64)
            rc = lib().spm_cuda_split_lines(self._h, buf.ctypes.data, len(buf), text.ctypes.data, off.ctypes.data,
                                            cap, ctypes.byref(n), HOST, stream)
            if rc == 8 and n.value > cap:      # RESOURCE_EXHAUSTED: n_lines holds the needed capacity
                cap = n.value
                continue
            _check(rc)
            off = off[: n.value + 1]
            return text[: int(off[-1])], off

    def split_lines_device(self, buf, text, offsets, noalloc=False, stream=None):
        """Device-resident line splitting on torch CUDA tensors (uint8 buf, uint8 text of >= buf's
        size, int64 offsets of capacity + 1); returns the line count.  noalloc=True takes
        spm_cuda_split_lines_noalloc: static device scratch, no allocation, buffers <= 256 MiB."""
        import torch
        self._need()
        if stream is None:
            stream = torch.cuda.current_stream(buf.device).cuda_stream
        n = ctypes.c_int64(0)
        cap = offsets.numel() - 1
        if noalloc:
            rc = lib().spm_cuda_split_lines_noalloc(self._h, buf.data_ptr(), buf.numel(), text.data_ptr(),
                                                    offsets.data_ptr(), cap, ctypes.byref(n), stream)
        else:
            rc = lib().spm_cuda_split_lines(self._h, buf.data_ptr(), buf.numel(), text.data_ptr(),
                                            offsets.data_ptr(), cap, ctypes.byref(n), DEVICE, stream)
        _check(rc)
        return n.value

    def encode_file_buffer(self, buf):
        """spm_encode --output_format=id on a raw file buffer: (ids int32[], id_offsets int64[n+1])."""
        text, off = self.split_lines(buf)
        return self.encode_packed(text, off)

    # ---- pieces / SentencePieceText (sentencepiece_processor.h:258, 338-381)
    def encode_pieces_packed(self, text, offsets, stream=None):
        """Host CSR in -> dict of host arrays: ids, begin, end (spans in the input), nbegin, nend
        (spans in the normalized text), id_off, norm (normalized text), norm_off
        (layout of spm_cuda_encode_batch_pieces, include/spm_cuda.h)."""
        self._need()
        text = np.ascontiguousarray(text, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        n = len(offsets) - 1
        raw = int(offsets[-1] - offsets[0])
        cap, ncap = max(64, raw // 2 + 4 * n), max(64, raw + raw // 2 + 8 * n)
        id_off = np.zeros(n + 1, dtype=np.int64)
        norm_off = np.zeros(n + 1, dtype=np.int64)
        while True:
            a = {k: np.empty(cap, dtype=np.int32) for k in ("ids", "begin", "end", "nbegin", "nend")}
            norm = np.empty(ncap, dtype=np.uint8)
            rc = lib().spm_cuda_encode_batch_pieces(
                self._h, text.ctypes.data, offsets.ctypes.data, n, self._opts, a["ids"].ctypes.data,
                a["begin"].ctypes.data, a["end"].ctypes.data, a["nbegin"].ctypes.data, a["nend"].ctypes.data, cap,
                id_off.ctypes.data, norm.ctypes.data, ncap, norm_off.ctypes.data, HOST, stream)
            if rc == 8 and (int(id_off[n]) > cap or int(norm_off[n]) > ncap):
                cap, ncap = max(cap, int(id_off[n])), max(ncap, int(norm_off[n]))
                continue
            _check(rc)
            out = {k: v[: id_off[n]] for k, v in a.items()}
            out.update(id_off=id_off, norm=norm[: norm_off[n]], norm_off=norm_off)
            return out

    def encode_pieces_device(self, text, offsets, ids, begin, end, nbegin, nend, id_offsets, norm_text, norm_offsets,
                             stream=None):
        """Device-resident variant on torch CUDA tensors (uint8, int64, 5 x int32 of equal size,
        int64, uint8, int64)."""
        import torch
        self._need()
        if stream is None:
            stream = torch.cuda.current_stream(text.device).cuda_stream
        n = offsets.numel() - 1
        _check(lib().spm_cuda_encode_batch_pieces(
            self._h, text.data_ptr(), offsets.data_ptr(), n, self._opts, ids.data_ptr(), begin.data_ptr(),
            end.data_ptr(), nbegin.data_ptr(), nend.data_ptr(), ids.numel(), id_offsets.data_ptr(),
            norm_text.data_ptr(), norm_text.numel(), norm_offsets.data_ptr(), DEVICE, stream))

    def _one(self, input):
        b = input.encode("utf-8") if isinstance(input, str) else bytes(input)
        text = np.frombuffer(b, dtype=np.uint8)
        off = np.array([0, len(b)], dtype=np.int64)
        return b, text, off, self.encode_pieces_packed(text, off)

    def EncodeAsPieces(self, input):
        b, text, off, res = self._one(input)
        return [p[0].decode("utf-8", "replace") for p in spt.sentence_pieces(0, text, off, res, self._piece_bytes)]

    def EncodeAsSerializedProto(self, input):
        b, text, off, res = self._one(input)
        return spt.serialize(b, spt.sentence_pieces(0, text, off, res, self._piece_bytes))

    def encode_as_serialized_protos(self, text, offsets):
        """Batch EncodeAsSerializedProto: list of serialized SentencePieceText, one per sentence."""
        text = np.ascontiguousarray(text, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        res = self.encode_pieces_packed(text, offsets)
        return [spt.serialize(text[offsets[i]:offsets[i + 1]].tobytes(),
                              spt.sentence_pieces(i, text, offsets, res, self._piece_bytes))
                for i in range(len(offsets) - 1)]

    # ---- decode (sentencepiece_processor.h:252, .cc:670-733)
    def decode_packed(self, ids, id_offsets, stream=None):
        """Host CSR ids -> (text uint8[], text_offsets int64[n+1])."""
        self._need()
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        id_offsets = np.ascontiguousarray(id_offsets, dtype=np.int64)
        n = len(id_offsets) - 1
        toff = np.zeros(n + 1, dtype=np.int64)
        cap = max(64, 6 * len(ids))
        while True:
            text = np.empty(cap, dtype=np.uint8)
            rc = lib().spm_cuda_decode_batch(self._h, ids.ctypes.data, id_offsets.ctypes.data, n, text.ctypes.data,
                                             cap, toff.ctypes.data, HOST, stream)
            if rc == 8 and int(toff[n]) > cap:
                cap = int(toff[n])
                continue
            _check(rc)
            return text[: toff[n]], toff

    def decode_device(self, ids, id_offsets, text, text_offsets, stream=None):
        """Device CSR in/out on torch CUDA tensors (int32, int64, uint8, int64)."""
        import torch
        self._need()
        if stream is None:
            stream = torch.cuda.current_stream(ids.device).cuda_stream
        n = id_offsets.numel() - 1
        _check(lib().spm_cuda_decode_batch(self._h, ids.data_ptr(), id_offsets.data_ptr(), n, text.data_ptr(),
                                           text.numel(), text_offsets.data_ptr(), DEVICE, stream))

    def DecodeIds(self, ids):
        text, _ = self.decode_packed(np.asarray(ids, dtype=np.int32), np.array([0, len(ids)], dtype=np.int64))
        return text.tobytes().decode("utf-8", "replace")

    Decode = DecodeIds

    # ---- dense / sparse layouts of the TF ops (tensorflow/sentencepiece_processor_ops.cc:330-428;
    # SURVEY 8(f) rank 3): torch CUDA tensors in and out
    def csr_to_dense(self, ids, id_offsets, pad=0, stream=None):
        """(ids int32[], id_offsets int64[n+1]) on the GPU -> (values int32[n, max_len], seq_len int32[n])."""
        import torch
        self._need()
        if stream is None:
            stream = torch.cuda.current_stream(ids.device).cuda_stream
        n = id_offsets.numel() - 1
        mx = ctypes.c_int64(0)
        _check(lib().spm_cuda_csr_max_len(self._h, id_offsets.data_ptr(), n, ctypes.byref(mx), stream))
        values = torch.empty((n, mx.value), dtype=torch.int32, device=ids.device)
        seq_len = torch.empty(n, dtype=torch.int32, device=ids.device)
        _check(lib().spm_cuda_csr_to_dense(self._h, ids.data_ptr(), id_offsets.data_ptr(), n, int(pad),
                                           values.data_ptr(), mx.value, seq_len.data_ptr(), stream))
        return values, seq_len

    def csr_to_sparse(self, ids, id_offsets, stream=None):
        """-> (indices int64[nnz, 2], values int32[nnz], dense_shape (n, max_len))."""
        import torch
        self._need()
        if stream is None:
            stream = torch.cuda.current_stream(ids.device).cuda_stream
        n = id_offsets.numel() - 1
        mx = ctypes.c_int64(0)
        _check(lib().spm_cuda_csr_max_len(self._h, id_offsets.data_ptr(), n, ctypes.byref(mx), stream))
        nnz = ids.numel()
        indices = torch.empty((nnz, 2), dtype=torch.int64, device=ids.device)
        _check(lib().spm_cuda_csr_to_sparse(self._h, id_offsets.data_ptr(), n, indices.data_ptr(), stream))
        return indices, ids, (n, mx.value)

    def encode_dense(self, sentences, pad=0):
        """The TF op SentencepieceEncodeDense on the GPU: list of str/bytes ->
        (values int32[n, max_len], seq_len int32[n]) as torch CUDA tensors."""
        import torch
        bs = [s.encode("utf-8") if isinstance(s, str) else bytes(s) for s in sentences]
        off = np.zeros(len(bs) + 1, dtype=np.int64)
        if bs:
            np.cumsum([len(b) for b in bs], out=off[1:])
        dev = torch.device("cuda", self._device)
        text = torch.from_numpy(np.frombuffer(b"".join(bs) + b"\0", dtype=np.uint8).copy()).to(dev)
        d_off = torch.from_numpy(off).to(dev)
        cap = int(off[-1]) + 4 * len(bs) + 64
        ids = torch.empty(cap, dtype=torch.int32, device=dev)
        id_off = torch.empty(len(bs) + 1, dtype=torch.int64, device=dev)
        self.encode_device(text, d_off, ids, id_off)
        total = int(id_off[-1].item())
        return self.csr_to_dense(ids[:total], id_off, pad)

    STAGES = ("encode", "normalize", "memo_collect", "memo_words", "memo_emit", "compact")

    def last_stats(self):
        """Counters of the last encode call (include/spm_cuda.h, spm_cuda_last_stats)."""
        s = (ctypes.c_double * 16)()
        lib().spm_cuda_last_stats(self._h, s, 16)
        return {"launches": int(s[0]), "fallback_sentences": int(s[1]), "device_ms": s[2],
                "workspace_bytes": int(s[3]), "encode_kernel_ms": s[4], "encode_kernel_launches": int(s[5]),
                "sentences": int(s[6]), "ids": int(s[7]), "dominant_stage": self.STAGES[int(s[8])],
                "stage_ms": {name: s[9 + k] for k, name in enumerate(self.STAGES)},
                "viterbi_tier_sentences": int(s[15])}

    # ---- model introspection (sentencepiece_processor.h:397-432)
    def GetPieceSize(self):
        self._need()
        return lib().spm_cuda_model_piece_size(self._h)

    def _piece_bytes(self, id):
        n = ctypes.c_int(0)
        p = lib().spm_cuda_model_piece(self._h, id, ctypes.byref(n))
        return ctypes.string_at(p, n.value)

    def IdToPiece(self, id):
        self._need()
        return self._piece_bytes(id).decode("utf-8", "replace")

    def PieceToId(self, piece):
        self._need()
        b = piece.encode("utf-8") if isinstance(piece, str) else bytes(piece)
        return lib().spm_cuda_model_piece_to_id(self._h, b, len(b))

    def GetScore(self, id):
        self._need()
        return lib().spm_cuda_model_score(self._h, id)

    def _type(self, id):
        self._need()
        return lib().spm_cuda_model_piece_type(self._h, id)

    def IsUnknown(self, id):
        return self._type(id) == UNKNOWN

    def IsControl(self, id):
        return self._type(id) == CONTROL

    def IsUnused(self, id):
        return self._type(id) == UNUSED

    def unk_id(self):
        self._need()
        return lib().spm_cuda_model_unk_id(self._h)

    def bos_id(self):
        self._need()
        return lib().spm_cuda_model_bos_id(self._h)

    def eos_id(self):
        self._need()
        return lib().spm_cuda_model_eos_id(self._h)

    def pad_id(self):
        self._need()
        return lib().spm_cuda_model_pad_id(self._h)

    def model_type(self):
        self._need()
        return lib().spm_cuda_model_type(self._h)

    # snake_case aliases of the reference's python module
    load = Load
    encode_as_ids = EncodeAsIds
    encode_as_pieces = EncodeAsPieces
    encode_as_serialized_proto = EncodeAsSerializedProto
    decode_ids = DecodeIds
    get_piece_size = GetPieceSize
    piece_to_id = PieceToId
    id_to_piece = IdToPiece
    __len__ = GetPieceSize


class UnigramEStep:
    """The E-step of the unigram trainer on one GPU (include/spm_cuda.h, spm_cuda_estep_*):
    what the worker threads of unigram::Trainer::RunEStep compute
    (src/unigram_model_trainer.cc:237-287) for the (word, freq) pairs given to run().
    The sum over threads that the reference does serially (:274-280) is, across GPUs, one
    all-reduce of the packed [expected | obj | ntokens] vector: sharding.allreduce_estep()."""

    def __init__(self, pieces, scores, device=0):
        import numpy as np
        self._h = None
        self._np = np
        ps = [p if isinstance(p, bytes) else p.encode("utf-8") for p in pieces]
        if any(b"\0" in p for p in ps):
            raise StatusError(3, "piece contains NUL")
        sc = np.ascontiguousarray(scores, dtype=np.float32)
        if len(sc) != len(ps):
            raise StatusError(3, "pieces and scores differ in length")
        blob = b"".join(p + b"\0" for p in ps)
        h = ctypes.c_void_p()
        _check(lib().spm_cuda_estep_create(blob, sc.ctypes.data, len(ps), device, ctypes.byref(h)))
        self._h = h
        self.vocab_size = len(ps)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().spm_cuda_estep_free(self._h)
            self._h = None

    def run(self, words, offsets, freq, all_freq=None, want_logz=False, stream=None):
        """words: uint8 array (all words back to back), offsets int64[n+1], freq int64[n] (host
        numpy arrays).  all_freq: sum of freq over ALL shards (default: this shard's).
        Returns (expected float64[V], obj, ntokens[, logz float32[n]])."""
        np = self._np
        words = np.ascontiguousarray(words, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        freq = np.ascontiguousarray(freq, dtype=np.int64)
        n = len(offsets) - 1
        if len(freq) != n:
            raise StatusError(3, "freq and offsets differ in length")
        if all_freq is None:
            all_freq = float(freq.sum())
        expected = np.zeros(self.vocab_size, dtype=np.float64)
        out = np.zeros(2, dtype=np.float64)
        logz = np.zeros(max(n, 1), dtype=np.float32) if want_logz else None
        _check(lib().spm_cuda_estep_run(self._h, words.ctypes.data, offsets.ctypes.data, freq.ctypes.data, n,
                                        float(all_freq), expected.ctypes.data, out.ctypes.data,
                                        logz.ctypes.data if want_logz else None, HOST, stream))
        res = (expected, float(out[0]), int(round(out[1])))
        return res + (logz[:n],) if want_logz else res

    def run_device(self, words, offsets, freq, all_freq, expected, obj_ntokens, stream=None):
        """Device-resident variant: torch CUDA tensors (uint8, int64, int64, float64[V], float64[2]);
        expected / obj_ntokens are accumulated into."""
        n = offsets.numel() - 1
        _check(lib().spm_cuda_estep_run(self._h, words.data_ptr(), offsets.data_ptr(), freq.data_ptr(), n,
                                        float(all_freq), expected.data_ptr(), obj_ntokens.data_ptr(), None,
                                        DEVICE, stream))
