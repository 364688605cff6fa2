# parity subset (memo / dirty / edge cases), then config 2 clean + dirty and config 3 dirty
V=${1:-vX}; O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -p no:cacheprovider -k "corpus_parity or edge_sentences or option_paths or host_chunk or test_smoke or memo or dirty or word_scan or 2m or small_memo" ) > $O/r2_${V}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2_${V}_pytest_gpu.log
grep -v "LOG(INFO)" $O/r2_${V}_pytest_gpu.log | tail -6
run() { local name=$1; shift
  timeout 600 python bench.py --no-configs "$@" > $O/r2_${V}_bench_$name.json 2> $O/r2_${V}_bench_$name.err; echo "bench $name rc=$?"; grep -v "LOG(INFO)" $O/r2_${V}_bench_$name.err | tail -3
  python - $O/r2_${V}_bench_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]/1e6,2), "M lines/s", round(d["ms_per_step"],3), "ms e2e", round(d.get("e2e",{}).get("ms_per_step",0),2), d.get("stage_ms_per_step"), (d.get("parity") or {}).get("sentences_mismatching"), d["viterbi_tier_sentences_per_step"], d["general_tier_sentences_per_step"])
except Exception as e: print("bad", e)
PY
}
run c2
run c2_dirty --corpus syn_dirty
run c3_dirty --model syn_bpe32k --corpus syn_dirty
