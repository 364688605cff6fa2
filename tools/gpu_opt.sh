# A/B of run-time options on one bench configuration: gpu_opt.sh vX "<bench args>" opt1=v[,opt2=v] ...
V=${1:-vX}; ARGS=$2; shift; shift; O=gpurun_out; mkdir -p $O
for spec in base "$@"; do
  OPTS=""; if [ "$spec" != base ]; then for kv in ${spec//,/ }; do OPTS="$OPTS --option $kv"; done; fi
  timeout 600 python bench.py --cpu-sample 60000 --no-configs --no-e2e $ARGS $OPTS > $O/r2_${V}_opt_${spec//[=,]/_}.json 2> $O/r2_${V}_opt.err; echo "bench $spec rc=$?"
  python - $O/r2_${V}_opt_${spec//[=,]/_}.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]/1e6,2), "M lines/s", round(d["ms_per_step"],3), "ms", d.get("stage_ms_per_step"), (d.get("parity") or {}).get("sentences_mismatching"))
except Exception as e: print("bad", e)
PY
done
