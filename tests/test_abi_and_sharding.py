"""CPU-side checks of the drop-in boundary and of the multi-GPU host logic:
  * libspm_cuda.so loads and exports every symbol include/spm_cuda.h declares;
  * without a GPU the product fails loudly (no CPU fallback);
  * sentence sharding (world_size 2, gloo) reproduces the single-process CSR."""
import ctypes
import os
import re
import socket

import numpy as np
import pytest

from conftest import ROOT, model_bytes


def declared_symbols():
    with open(os.path.join(ROOT, "include", "spm_cuda.h")) as f:
        src = f.read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(spm_cuda_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_declared_symbols(spm_b200):
    L = spm_b200.lib()
    names = declared_symbols()
    assert len(names) >= 15
    for name in names:
        assert hasattr(L, name), "libspm_cuda.so does not export %s" % name


def test_pack_extra_options(spm_b200):
    assert spm_b200.pack_extra_options("") == 0
    assert spm_b200.pack_extra_options("bos:eos:reverse") == 1 | (2 << 2) | (3 << 4)
    with pytest.raises(spm_b200.StatusError):
        spm_b200.pack_extra_options("foo")


def test_no_cpu_fallback(spm_b200):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    sp = spm_b200.SentencePieceProcessor()
    with pytest.raises(spm_b200.StatusError) as e:
        sp.LoadFromSerializedProto(model_bytes("botchan_uni8k"))
    assert "no CUDA device" in str(e.value)
    with pytest.raises(spm_b200.StatusError):
        sp.EncodeAsIds("hello")


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "sentencepiece-comments_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cc", ".h")):
                with open(os.path.join(dirpath, fn), errors="replace") as f:
                    src = f.read()
                assert not re.search(r"^\s*(from|import)\s+oracle|#include\s+\"[^\"]*oracle|_oracle\.so|libspm_ref", src,
                                     flags=re.M), "%s reaches into oracle/" % fn


def test_shard_ranges(spm_b200, corpora):
    from spm_b200 import sharding
    _, off = corpora["botchan"]
    for world in (1, 2, 3, 8):
        r = sharding.shard_ranges(off, world)
        assert r[0][0] == 0 and r[-1][1] == len(off) - 1
        assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
        sizes = [int(off[hi] - off[lo]) for lo, hi in r]
        assert max(sizes) - min(sizes) <= 2 * int(np.diff(off).max())
    assert sharding.shard_ranges(np.zeros(1, dtype=np.int64), 4) == [(0, 0)] * 4


def _worker(rank, world, port_no, q):
    import torch.distributed as dist
    import spm_b200_pkg
    from oracle import port
    from tools import synth
    pkg = spm_b200_pkg.load()
    from spm_b200 import sharding
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port_no, rank=rank, world_size=world)
    text, off = synth.pack_lines(synth.read_gz_lines("corpus_en_botchan.txt.gz")[:1200])
    ltext, loff, (lo, hi) = sharding.local_shard(text, off, rank, world)
    # the CPU stand-in for this rank's GPU encode (the oracle is only the checker here)
    ids, io = port.OracleProcessor(model_bytes("botchan_uni8k")).encode_batch(ltext, loff)
    gio, total = sharding.global_id_offsets(io)
    q.put((rank, lo, hi, ids, gio, total))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_encode_world2_gloo():
    import torch.multiprocessing as mp
    from oracle import port
    from tools import synth
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port_no = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    world = 2
    procs = [ctx.Process(target=_worker, args=(r, world, port_no, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=300) for _ in range(world)], key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    text, off = synth.pack_lines(synth.read_gz_lines("corpus_en_botchan.txt.gz")[:1200])
    wi, wo = port.OracleProcessor(model_bytes("botchan_uni8k")).encode_batch(text, off)
    assert got[0][1] == 0 and got[0][2] == got[1][1] and got[1][2] == 1200
    assert got[0][5] == got[1][5] == len(wi)
    ids = np.concatenate([g[3] for g in got])
    offs = np.concatenate([got[0][4][:-1], got[1][4]])
    assert np.array_equal(ids, wi) and np.array_equal(offs, wo)


def _estep_worker(rank, world, port_no, q):
    import torch.distributed as dist
    import spm_b200_pkg
    from oracle import ref
    spm_b200_pkg.load()
    from spm_b200 import sharding
    import test_estep as te
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port_no, rank=rank, world_size=world)
    pieces, scores, words, freq = te.trainer_like_case(1500)
    _, off = te.pack_words(words)
    lo, hi = sharding.estep_shard_ranges(off, freq, world)[rank]
    # the CPU stand-in for this rank's GPU E-step (the reference is only the checker here)
    exp, _, obj, nt = ref.estep(pieces, scores, words[lo:hi], freq[lo:hi])
    obj *= float(freq[lo:hi].sum()) / float(freq.sum())     # ref_estep divides by ITS sum of freq
    exp, obj, nt = sharding.allreduce_estep(exp.astype(np.float64), obj, nt)
    q.put((rank, exp, obj, nt))
    dist.barrier()
    dist.destroy_process_group()


def test_estep_allreduce_world2_gloo():
    """The one exchange step of the path: expected[V] | obj | ntokens summed over the ranks."""
    import torch.multiprocessing as mp
    from oracle import ref
    import test_estep as te
    if not ref.available() or not hasattr(ref.lib(), "ref_estep"):
        pytest.skip("reference library with the trainer is not built")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port_no = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    world = 2
    procs = [ctx.Process(target=_estep_worker, args=(r, world, port_no, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=300) for _ in range(world)], key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    pieces, scores, words, freq = te.trainer_like_case(1500)
    wexp, _, wobj, wnt = ref.estep(pieces, scores, words, freq)
    for _, exp, obj, nt in got:
        assert nt == wnt
        assert np.allclose(exp, wexp, rtol=1e-4, atol=1e-5)
        assert abs(obj - wobj) <= 1e-4 * abs(wobj)
    assert np.array_equal(got[0][1], got[1][1])
