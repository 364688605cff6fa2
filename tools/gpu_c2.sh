# config 2 (clean, dirty), config 3 and the launch list of config 2
V=${1:-vX}; O=gpurun_out; mkdir -p $O
run() { local name=$1; shift
  timeout 600 python bench.py --no-configs "$@" > $O/r2_${V}_bench_$name.json 2> $O/r2_${V}_bench_$name.err; echo "bench $name rc=$?"; grep -v "LOG(INFO)" $O/r2_${V}_bench_$name.err | tail -3
  python - $O/r2_${V}_bench_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]/1e6,2), "M lines/s", round(d["ms_per_step"],3), "ms", round(d["input_gb_per_s"],2),"GB/s e2e", round(d.get("e2e",{}).get("ms_per_step",0),2), d.get("stage_ms_per_step"), (d.get("parity") or {}).get("sentences_mismatching"), d["viterbi_tier_sentences_per_step"], d["general_tier_sentences_per_step"], "frac", round(d["roofline"]["frac"],4))
except Exception as e: print("bad", e)
PY
}
run c2
run c2_dirty --corpus syn_dirty
run c3 --model syn_bpe32k
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_${V}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-configs > $O/r2_${V}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
