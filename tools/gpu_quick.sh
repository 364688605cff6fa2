# parity subset + config 2 (clean, dirty), config 3, config 4
V=${1:-vX}; O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -p no:cacheprovider -k "corpus_parity or edge_sentences or option_paths or long_lines or encode_fast_kernel or test_smoke or 2m" ) > $O/r2_${V}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2_${V}_pytest_gpu.log
grep -v "LOG(INFO)" $O/r2_${V}_pytest_gpu.log | tail -4
for cfg in "c2:" "dirty:--corpus syn_dirty" "c3:--model syn_bpe32k" "c4:--model ref_ja_8k --corpus wagahai --sentences 2200000" "c2_nomemo:--option memo=0 --sentences 2000000"; do
  tag=${cfg%%:*}; args=${cfg#*:}
  timeout 600 python bench.py --cpu-sample 60000 --no-configs --no-e2e $args > $O/r2_${V}_q_$tag.json 2> $O/r2_${V}_q.err; echo "bench $tag rc=$?"
  python - $O/r2_${V}_q_$tag.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["ms_per_step"],3), "ms", d.get("stage_ms_per_step"), (d.get("parity") or {}).get("sentences_mismatching"))
except Exception as e: print("bad", e)
PY
done
