V=${1:-vX}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x --timeout 900 -p no:cacheprovider -k "corpus_parity or dirty_1m or option_paths or edge_sentences" 2>&1 | grep -v LOG | tail -3
timeout 600 python bench.py --cpu-sample 200000 --no-e2e --no-configs --corpus syn_dirty > $O/r2_${V}_bench_config2_dirty.json 2>/dev/null
shopt -s nullglob
for f in sentencepiece-comments_b200/lib/variants/*.so; do
  n=$(basename $f .so)
  SPM_B200_LIB_VARIANT=$PWD/$f timeout 300 python bench.py --cpu-sample 200000 --no-e2e --no-configs --corpus syn_dirty > $O/r2_${V}_bench_config2_dirty_var_$n.json 2>/dev/null
done
for f in $O/r2_${V}_bench_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]/1e6,1), "M/s", round(d["ms_per_step"],3), "ms", d.get("stage_ms_per_step"), (d.get("parity") or {}).get("sentences_mismatching"))
except Exception as e: print("bad", e)
PY
done
