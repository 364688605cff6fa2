"""Minimal proto2 wire writer/reader for sentencepiece ModelProto (test helper).

Field numbers follow /root/reference/src/sentencepiece_model.proto.
Used to build the small models of the reference's own unit tests
(unigram_model_test.cc, bpe_model_test.cc, sentencepiece_processor_test.cc,
normalizer_test.cc) without a protobuf runtime.
"""
import struct

NORMAL, UNKNOWN, CONTROL, USER_DEFINED, UNUSED = 1, 2, 3, 4, 5
UNIGRAM, BPE, WORD, CHAR = 1, 2, 3, 4


def _varint(v):
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _key(num, wt):
    return _varint((num << 3) | wt)


def _ld(num, payload):
    return _key(num, 2) + _varint(len(payload)) + payload


def _vi(num, v):
    return _key(num, 0) + _varint(int(v))


def piece(text, score=0.0, type=NORMAL):
    b = text if isinstance(text, bytes) else text.encode("utf-8")
    msg = _ld(1, b) + _key(2, 5) + struct.pack("<f", score)
    if type != NORMAL:
        msg += _vi(3, type)
    return msg


def model(pieces, model_type=UNIGRAM, charsmap=b"", add_dummy_prefix=True, remove_extra_whitespaces=True,
          escape_whitespaces=True, treat_whitespace_as_suffix=False):
    """pieces: list of (text, score[, type]) -> serialized ModelProto."""
    out = b""
    for p in pieces:
        out += _ld(1, piece(*p))
    ts = _vi(3, model_type)
    if treat_whitespace_as_suffix:
        ts += _vi(24, 1)
    out += _ld(2, ts)
    ns = b""
    if charsmap:
        ns += _ld(2, charsmap)
    ns += _vi(3, add_dummy_prefix) + _vi(4, remove_extra_whitespaces) + _vi(5, escape_whitespaces)
    out += _ld(3, ns)
    return out


def base_pieces():
    """MakeBaseModelProto of the reference tests: <unk>, <s>, </s>."""
    return [("<unk>", 0.0, UNKNOWN), ("<s>", 0.0, CONTROL), ("</s>", 0.0, CONTROL)]


def _fields(buf):
    i, n = 0, len(buf)
    while i < n:
        k = 0
        sh = 0
        while True:
            b = buf[i]
            i += 1
            k |= (b & 0x7F) << sh
            sh += 7
            if not b & 0x80:
                break
        num, wt = k >> 3, k & 7
        if wt == 0:
            v = 0
            sh = 0
            while True:
                b = buf[i]
                i += 1
                v |= (b & 0x7F) << sh
                sh += 7
                if not b & 0x80:
                    break
            yield num, wt, v
        elif wt == 1:
            yield num, wt, buf[i:i + 8]
            i += 8
        elif wt == 5:
            yield num, wt, buf[i:i + 4]
            i += 4
        elif wt == 2:
            ln = 0
            sh = 0
            while True:
                b = buf[i]
                i += 1
                ln |= (b & 0x7F) << sh
                sh += 7
                if not b & 0x80:
                    break
            yield num, wt, buf[i:i + ln]
            i += ln
        else:
            raise ValueError("unsupported wire type %d" % wt)


def charsmap_of(model_bytes):
    """precompiled_charsmap (NormalizerSpec field 2) of a serialized model."""
    for num, wt, v in _fields(model_bytes):
        if num == 3 and wt == 2:
            for n2, w2, v2 in _fields(v):
                if n2 == 2 and w2 == 2:
                    return bytes(v2)
    return b""


def pieces_of(model_bytes):
    """[(piece bytes, score, type)] of a serialized model."""
    out = []
    for num, wt, v in _fields(model_bytes):
        if num == 1 and wt == 2:
            p, sc, ty = b"", 0.0, NORMAL
            for n2, w2, v2 in _fields(v):
                if n2 == 1:
                    p = bytes(v2)
                elif n2 == 2:
                    sc = struct.unpack("<f", v2)[0]
                elif n2 == 3:
                    ty = v2
            out.append((p, sc, ty))
    return out
