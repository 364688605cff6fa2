# A/B of build variants on config 4 (2.2 M lines): stage times per variant
V=${1:-vX}; shift; O=gpurun_out; mkdir -p $O
for name in base "$@"; do
  if [ "$name" = base ]; then unset SPM_B200_LIB_VARIANT; else export SPM_B200_LIB_VARIANT=$PWD/sentencepiece-comments_b200/lib/variants/libspm_cuda_$name.so; fi
  timeout 600 python bench.py --cpu-sample 60000 --no-configs --no-e2e $BENCH_ARGS > $O/r2_${V}_var_$name.json 2> $O/r2_${V}_var_$name.err; echo "bench $name rc=$?"
  python - $O/r2_${V}_var_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]/1e6,2), "M lines/s", round(d["ms_per_step"],3), "ms", d.get("stage_ms_per_step"), (d.get("parity") or {}).get("sentences_mismatching"))
except Exception as e: print("bad", e)
PY
done
