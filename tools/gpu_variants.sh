# GPU tests + bench of the default build, then bench lines of differently compiled variants
# (sentencepiece-comments_b200/lib/variants/*.so, loaded through SPM_B200_LIB_VARIANT), one ncu capture.
# Usage: gpurun --timeout 1200 -- 'bash tools/gpu_variants.sh v21'
V=${1:-vX}
K=${2:-k_memo_scan_words|k_memo_emit|k_compact}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${V}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${V}_pytest_gpu.log; tail -3 $O/${V}_pytest_gpu.log
timeout 300 python bench.py --no-cpu-baseline > $O/${V}_bench_config2.json 2> $O/${V}_bench_config2.err; echo "bench rc=$?"
timeout 300 python bench.py --corpus syn_dirty --no-cpu-baseline --no-e2e > $O/${V}_bench_config2_dirty.json 2>/dev/null
timeout 300 python bench.py --model syn_bpe32k --no-cpu-baseline --no-e2e > $O/${V}_bench_config3.json 2>/dev/null
shopt -s nullglob
for f in sentencepiece-comments_b200/lib/variants/*.so; do
  n=$(basename $f .so)
  SPM_B200_LIB_VARIANT=$PWD/$f timeout 300 python bench.py --no-cpu-baseline --no-e2e > $O/${V}_bench_config2_var_$n.json 2>/dev/null
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$K" -c 3 -f -o $O/${V}_full_ids python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $O/${V}_ncu_full_ids.log 2>&1
for f in $O/${V}_bench_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(d["value"], d["ms_per_step"], d.get("e2e",{}).get("value"), d.get("stage_ms_per_step"))
except Exception as e: print("bad", e)
PY
done
