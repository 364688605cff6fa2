# pieces path: parity (GPU pieces tests), then the pieces bench on config 2 (clean, dirty), config 3 and config 4, A/B of pieces_plain
V=${1:-vX}; O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -p no:cacheprovider -k "pieces or facade or test_smoke" ) > $O/r2_${V}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2_${V}_pytest_gpu.log
grep -v "LOG(INFO)" $O/r2_${V}_pytest_gpu.log | tail -6
run() { local name=$1; shift
  timeout 600 python tools/pieces_bench.py --op pieces "$@" > $O/r2_${V}_pieces_$name.json 2> $O/r2_${V}_pieces_$name.err; echo "pieces $name rc=$?"; grep -v "LOG(INFO)" $O/r2_${V}_pieces_$name.err | tail -2
  python - $O/r2_${V}_pieces_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]/1e6,2), "M/s", round(d["ms_per_step"],3), "ms e2e", round(d.get("e2e",{}).get("ms_per_step",0),2), d.get("stage_ms_per_step"), d.get("parity"))
except Exception as e: print("bad", e)
PY
}
run c2 --sentences 10000000
run c2_dirty --sentences 10000000 --corpus syn_dirty
run c3 --sentences 10000000 --model syn_bpe32k
run c4 --sentences 200000 --model ref_ja_8k --corpus wagahai
