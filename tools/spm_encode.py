"""spm_encode --output_format=id on a B200 (src/spm_encode_main.cc:52-204).

    python tools/spm_encode.py --model m.model [--extra_options bos:eos] [--output out.txt] input.txt ...

Lines are split on the GPU (spm_cuda_split_lines: std::getline semantics), encoded in one
batch call per file and written as space-separated ids, one line per input line."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def format_ids(ids, off):
    """'1 2 3\\n4 5\\n' for CSR (ids, off) without a Python loop over the ids."""
    n = len(off) - 1
    if n == 0:
        return b""
    toks = np.char.mod("%d", ids).astype("S") if len(ids) else np.zeros(0, dtype="S1")
    seps = np.full(len(ids), b" ", dtype="S1")
    last = off[1:][np.diff(off) > 0] - 1
    seps[last] = b"\n"
    body = np.char.add(toks, seps)
    # empty lines have no ids: their newline goes in front of the next id (or at the end)
    empty_before = np.zeros(len(ids) + 1, dtype=np.int64)
    np.add.at(empty_before, off[:-1][np.diff(off) == 0], 1)
    out = []
    pos = 0
    for k in np.nonzero(empty_before)[0]:
        out.append(b"".join(body[pos:k].tolist()))
        out.append(b"\n" * int(empty_before[k]))
        pos = k
    out.append(b"".join(body[pos:].tolist()))
    return b"".join(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", required=True)
    ap.add_argument("--output", default="")
    ap.add_argument("--output_format", default="id", choices=["id"])
    ap.add_argument("--extra_options", default="")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("inputs", nargs="+")
    args = ap.parse_args()
    import spm_b200_pkg
    spm = spm_b200_pkg.load()
    sp = spm.SentencePieceProcessor(model_file=args.model, device=args.device)
    if args.extra_options:
        sp.SetEncodeExtraOptions(args.extra_options)
    out = open(args.output, "wb") if args.output else sys.stdout.buffer
    for name in args.inputs:
        with open(name, "rb") as f:
            ids, off = sp.encode_file_buffer(f.read())
        out.write(format_ids(ids, off))
    if args.output:
        out.close()


if __name__ == "__main__":
    main()
