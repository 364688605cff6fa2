# Config-4 pass: parity subset, then config 4 at 20 k / 200 k / 2.2 M lines and config 2 (clean, dirty).
V=${1:-vX}; O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -p no:cacheprovider -k "corpus_parity or long_lines or near_tie or encode_fast_kernel or edge_sentences or option_paths or host_chunk or test_smoke" ) > $O/r2_${V}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2_${V}_pytest_gpu.log
grep -v "LOG(INFO)" $O/r2_${V}_pytest_gpu.log | tail -25
run() { # name args...
  local name=$1; shift
  timeout 600 python bench.py --cpu-sample 60000 --no-configs "$@" > $O/r2_${V}_bench_$name.json 2> $O/r2_${V}_bench_$name.err; echo "bench $name rc=$?"; grep -v "LOG(INFO)" $O/r2_${V}_bench_$name.err | tail -3
}
run c4_200k --model ref_ja_8k --corpus wagahai --sentences 200000
run c4_200k_noroute --model ref_ja_8k --corpus wagahai --sentences 200000 --option warp_route=0
run c4_20k --model ref_ja_8k --corpus wagahai --sentences 20000
run c4_2200k --model ref_ja_8k --corpus wagahai --sentences 2200000
run c4_wuni --model wagahai_uni --corpus wagahai --sentences 2200000
run c2_10m
run c2_dirty --corpus syn_dirty
for f in $O/r2_${V}_bench_c*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]/1e6,2), "M lines/s", round(d["ms_per_step"],3), "ms", round(d["input_gb_per_s"],2),"GB/s e2e", round(d.get("e2e",{}).get("ms_per_step",0),2), d.get("stage_ms_per_step"), (d.get("parity") or {}).get("sentences_mismatching"), d["viterbi_tier_sentences_per_step"], d["general_tier_sentences_per_step"])
except Exception as e: print("bad", e)
PY
done
