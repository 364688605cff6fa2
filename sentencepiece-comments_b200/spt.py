"""SentencePieceText (src/sentencepiece.proto) from the CSR arrays of
spm_cuda_encode_batch_pieces: the piece / surface strings and the serialized proto that
SentencePieceProcessor::EncodeAsPieces / EncodeAsSerializedProto return
(src/sentencepiece_processor.h:338-381, .cc:488-551,752-758).  Host-side formatting only;
every number in it comes from the GPU path."""


def _varint(v):
    out = bytearray()
    v &= (1 << 64) - 1
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _ld(num, payload):
    return _varint((num << 3) | 2) + _varint(len(payload)) + payload


def _vi(num, v):
    return _varint(num << 3) + _varint(v)


def sentence_pieces(i, text, off, res, id_to_piece):
    """Pieces of sentence i as tuples (piece bytes, id, surface bytes or None, begin, end);
    surface None marks a bos / eos piece of the extra options (no span in the reference)."""
    t0 = int(off[i])
    n0 = int(res["norm_off"][i])
    # the reference takes the surface from a string_view built over input.data() as a C string
    # (absl::ClippedSubstr(input.data(), ...), sentencepiece_processor.cc:519-520): it ends at the
    # first NUL byte of the sentence; begin / end are not clipped
    sent = bytes(text[t0:int(off[i + 1])])
    clip = sent.find(b"\0")
    clip = len(sent) if clip < 0 else clip
    out = []
    for k in range(int(res["id_off"][i]), int(res["id_off"][i + 1])):
        pid = int(res["ids"][k])
        nb, ne = int(res["nbegin"][k]), int(res["nend"][k])
        if nb < 0:
            out.append((id_to_piece(pid), pid, None, 0, 0))
            continue
        b, e = int(res["begin"][k]), int(res["end"][k])
        out.append((bytes(res["norm"][n0 + nb:n0 + ne]), pid, sent[min(b, clip):min(e, clip)], b, e))
    return out


def serialize(text_bytes, pieces):
    """SentencePieceText::SerializeAsString: text = 1, pieces = 2 {piece = 1, id = 2,
    surface = 3, begin = 4, end = 5}; proto2 presence: a bos / eos piece only has piece and id."""
    out = [_ld(1, bytes(text_bytes))]
    for piece, pid, surface, b, e in pieces:
        m = _ld(1, piece) + _vi(2, pid)
        if surface is not None:
            m += _ld(3, surface) + _vi(4, b) + _vi(5, e)
        out.append(_ld(2, m))
    return b"".join(out)
