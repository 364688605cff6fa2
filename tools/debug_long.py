import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from tools import synth
import spm_b200_pkg
from conftest import model_bytes, checker
text, off = synth.pack_lines(synth.read_gz_lines("corpus_en_botchan.txt.gz"))
reps = -(-(2 << 20) // len(text))
off = np.concatenate([[0]] + [off[1:] + r * len(text) for r in range(reps)]).astype(np.int64)
text = np.tile(text, reps)
rng = np.random.default_rng(5)
n = len(off) - 1
cuts = [0]
targets = [4095, 4096, 4097, 5000, 9000, 20000, 33000, 65535, 65536, 70000, 150000]
k = 0; i = 0
while i < n:
    if k < 200 and rng.random() < 0.05:
        want = targets[k % len(targets)] if k < 2 * len(targets) else int(rng.integers(4096, 40000))
        k += 1
        end = min(int(off[i]) + want, int(off[n]))
        while end < int(off[n]) and (text[end] & 0xC0) == 0x80: end += 1
        cuts.append(end)
        i = int(np.searchsorted(off, end, side="left"))
        if i <= n and i < len(off) and off[i] != end: cuts.append(int(off[i]))
    else:
        i += 1; cuts.append(int(off[i]))
cuts = np.unique(np.asarray(cuts, dtype=np.int64))
lens = np.diff(cuts)
pkg = spm_b200_pkg.load()
mb = model_bytes("botchan_uni8k")
sp = pkg.SentencePieceProcessor(); sp.LoadFromSerializedProto(mb)
chk, _ = checker(mb)
wi, wo = chk.encode_batch(text, cuts, threads=8)
for rep in range(3):
    gi, go = sp.encode_packed(text, cuts)
    bad = np.nonzero(np.diff(go) != np.diff(wo))[0]
    print("route", os.environ.get("SPM_CUDA_WARP_ROUTE"), "rep", rep, "bad sentences", [(int(b), int(lens[b]), int(np.diff(go)[b]), int(np.diff(wo)[b])) for b in bad[:20]], sp.last_stats()["viterbi_tier_sentences"], sp.last_stats()["fallback_sentences"])
