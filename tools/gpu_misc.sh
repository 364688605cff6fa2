V=${1:-vX}; O=gpurun_out; mkdir -p $O
for c in 12 16 20 24; do
timeout 300 python bench.py --no-cpu-baseline --no-configs --option host_chunks=$c > $O/r2_${V}_bench_chunks$c.json 2>/dev/null
done
for c in 5 8 12; do
timeout 300 python bench.py --no-cpu-baseline --no-configs --model syn_bpe32k --option host_chunks=$c > $O/r2_${V}_bench_bpe_chunks$c.json 2>/dev/null
done
for f in $O/r2_${V}_bench_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]/1e6,1), "M/s", round(d["ms_per_step"],3), "ms e2e", round(d.get("e2e",{}).get("ms_per_step",0),2))
except Exception as e: print("bad", e)
PY
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_memo_scan$' -c 1 -f -o $O/r2_${V}_full_memo_scan python bench.py --corpus syn_dirty --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-configs > $O/r2_${V}_ncu1.log 2>&1; echo "ncu rc=$?"
