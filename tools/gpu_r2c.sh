V=${1:-vX}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_estep.py -m gpu -q -x -p no:cacheprovider 2>&1 | grep -v LOG | tail -5
timeout 600 python tools/estep_bench.py > $O/r2_${V}_estep_1gpu.json 2> $O/r2_${V}_estep.err; echo "estep rc=$?"; grep -v LOG $O/r2_${V}_estep.err | tail -5
python - $O/r2_${V}_estep_1gpu.json <<'PY'
import json,sys
v=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
for n in ("final","seed"): print(n, round(v[n]["value"]/1e6,1),"Mwords/s", round(v[n]["ms_per_step"],3), v[n].get("parity"), v[n].get("cpu_baseline"))
PY
timeout 600 python bench.py --cpu-sample 500000 --no-configs > $O/r2_${V}_bench_config2.json 2> $O/r2_${V}_bench_config2.err; echo "bench rc=$?"; grep -v "LOG(INFO)" $O/r2_${V}_bench_config2.err | tail -5
shopt -s nullglob
for f in sentencepiece-comments_b200/lib/variants/*.so; do
  n=$(basename $f .so)
  SPM_B200_LIB_VARIANT=$PWD/$f timeout 300 python bench.py --cpu-sample 200000 --no-e2e --no-configs > $O/r2_${V}_bench_config2_var_$n.json 2>/dev/null
done
for f in $O/r2_${V}_bench_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]/1e6,1), "M/s", round(d["ms_per_step"],3), "ms e2e", round(d.get("e2e",{}).get("ms_per_step",0),2), d.get("stage_ms_per_step"), (d.get("parity") or {}).get("sentences_mismatching"), d["viterbi_tier_sentences_per_step"])
except Exception as e: print("bad", e)
PY
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_estep$' -c 2 -f -o $O/r2_${V}_full_estep python tools/estep_bench.py --no-reference --steps 1 --warmup 1 > $O/r2_${V}_ncu_estep.log 2>&1; echo "ncu rc=$?"
