# A/B of a build variant on config 2 (clean, dirty) and config 4
V=${1:-vX}; NAME=$2; O=gpurun_out; mkdir -p $O
for name in base $NAME; do
  if [ "$name" = base ]; then unset SPM_B200_LIB_VARIANT; else export SPM_B200_LIB_VARIANT=$PWD/sentencepiece-comments_b200/lib/variants/libspm_cuda_$name.so; fi
  for cfg in "c2:" "dirty:--corpus syn_dirty" "c4:--model ref_ja_8k --corpus wagahai --sentences 2200000"; do
    tag=${cfg%%:*}; args=${cfg#*:}
    timeout 600 python bench.py --cpu-sample 60000 --no-configs --no-e2e $args > $O/r2_${V}_var2_${name}_$tag.json 2> $O/r2_${V}_var2.err; echo "bench $name $tag rc=$?"
    python - $O/r2_${V}_var2_${name}_$tag.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["ms_per_step"],3), "ms", d.get("stage_ms_per_step"), (d.get("parity") or {}).get("sentences_mismatching"))
except Exception as e: print("bad", e)
PY
  done
done
