V=${1:-vX}; O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_estep.py -m gpu -q -x -p no:cacheprovider 2>&1 | grep -v LOG | tail -3
timeout 600 python tools/estep_bench.py > $O/r2_${V}_estep_1gpu.json 2> $O/r2_${V}_estep.err; echo "estep rc=$?"; grep -v LOG $O/r2_${V}_estep.err | tail -5
python - $O/r2_${V}_estep_1gpu.json <<'PY'
import json,sys
v=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
for n in ("final","seed"): print(n, round(v[n]["value"]/1e6,1),"Mwords/s", round(v[n]["ms_per_step"],3), v[n]["parity"]["ok"], v[n]["parity"]["max_rel_err_expected_vs_f64_reaccumulation"])
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'^k_estep$' -c 1 -f -o $O/r2_${V}_full_estep python tools/estep_bench.py --no-reference --vocabs final --steps 1 --warmup 0 > $O/r2_${V}_ncu_estep.log 2>&1; echo "ncu estep rc=$?"; tail -2 $O/r2_${V}_ncu_estep.log
