# Perf iteration pass: parity subset, config-2 line (default build, option variants, A/B library variants).
V=${1:-vX}; O=gpurun_out; mkdir -p $O
( time timeout 1200 python -m pytest tests -m gpu -q -x --timeout 900 -p no:cacheprovider -k "corpus_parity or word_scan or config_2m or dirty_1m or host_chunk or small_memo or option_paths or near_tie or extra_options or long_lines or memo_overflow" ) > $O/r2_${V}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2_${V}_pytest_gpu.log
grep -v "LOG(INFO)" $O/r2_${V}_pytest_gpu.log | tail -25
timeout 600 python bench.py --cpu-sample 500000 --no-configs > $O/r2_${V}_bench_config2.json 2> $O/r2_${V}_bench_config2.err; echo "bench rc=$?"; grep -v "LOG(INFO)" $O/r2_${V}_bench_config2.err | tail -5
timeout 600 python bench.py --cpu-sample 200000 --no-e2e --no-configs --option memo_bits=22 > $O/r2_${V}_bench_config2_bits22.json 2>/dev/null
shopt -s nullglob
for f in sentencepiece-comments_b200/lib/variants/*.so; do
  n=$(basename $f .so)
  SPM_B200_LIB_VARIANT=$PWD/$f timeout 300 python bench.py --cpu-sample 200000 --no-e2e --no-configs > $O/r2_${V}_bench_config2_var_$n.json 2>/dev/null
done
for f in $O/r2_${V}_bench_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]/1e6,1), "M/s", round(d["ms_per_step"],3), "ms e2e", round(d.get("e2e",{}).get("ms_per_step",0),2), d.get("stage_ms_per_step"), (d.get("parity") or {}).get("sentences_mismatching"), d["viterbi_tier_sentences_per_step"], d["general_tier_sentences_per_step"])
except Exception as e: print("bad", e)
PY
done
