# Config-4 pass: parity subset (long lines, Japanese corpora, near ties), then the config-4 bench lines at 2.2 M and 200 k lines.
V=${1:-vX}; O=gpurun_out; mkdir -p $O
( time timeout 1200 python -m pytest tests -m gpu -q -x --timeout 900 -p no:cacheprovider -k "corpus_parity or long_lines or near_tie or encode_fast_kernel or edge_sentences or option_paths or host_chunk" ) > $O/r2_${V}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2_${V}_pytest_gpu.log
grep -v "LOG(INFO)" $O/r2_${V}_pytest_gpu.log | tail -25
for n in 2200000 200000; do
timeout 600 python bench.py --model ref_ja_8k --corpus wagahai --sentences $n --cpu-sample 60000 --no-configs > $O/r2_${V}_bench_config4_$n.json 2> $O/r2_${V}_bench_config4.err; echo "bench rc=$?"; grep -v "LOG(INFO)" $O/r2_${V}_bench_config4.err | tail -3
done
timeout 600 python bench.py --model wagahai_uni --corpus wagahai --sentences 2200000 --cpu-sample 60000 --no-configs > $O/r2_${V}_bench_config4_wagahai_uni.json 2>/dev/null
for f in $O/r2_${V}_bench_config4*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]/1e6,2), "M lines/s", round(d["ms_per_step"],3), "ms", round(d["input_gb_per_s"],2),"GB/s e2e", round(d.get("e2e",{}).get("ms_per_step",0),2), d.get("stage_ms_per_step"), (d.get("parity") or {}).get("sentences_mismatching"), d["viterbi_tier_sentences_per_step"], d["general_tier_sentences_per_step"])
except Exception as e: print("bad", e)
PY
done
