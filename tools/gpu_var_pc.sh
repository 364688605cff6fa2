# A/B of build variants on the pieces path (config 2, 10 M sentences)
V=${1:-vX}; shift; O=gpurun_out; mkdir -p $O
for name in base "$@"; do
  if [ "$name" = base ]; then unset SPM_B200_LIB_VARIANT; else export SPM_B200_LIB_VARIANT=$PWD/sentencepiece-comments_b200/lib/variants/libspm_cuda_$name.so; fi
  timeout 600 python tools/pieces_bench.py --op pieces --sentences 10000000 --no-cpu-baseline --no-e2e > $O/r2_${V}_pcvar_$name.json 2> $O/r2_${V}_pcvar_$name.err; echo "bench $name rc=$?"
  python - $O/r2_${V}_pcvar_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]/1e6,2), "M/s", round(d["ms_per_step"],3), "ms", d.get("stage_ms_per_step"))
except Exception as e: print("bad", e)
PY
done
