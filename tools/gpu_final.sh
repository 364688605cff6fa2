# Final pass of a round: all GPU tests, smoke, the driver-shaped bench line (+ reference arm), launch list, ncu --set full of the
# config-2 kernels (-> profiles/traffic.json is refreshed from it by tools/ncu_traffic.py in the build container).
V=${1:-vX}; O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -p no:cacheprovider ) > $O/r2_${V}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2_${V}_pytest_gpu.log
grep -v "LOG(INFO)" $O/r2_${V}_pytest_gpu.log | tail -12
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -1
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep -v LOG | tail -4
( time timeout 900 python bench.py ) > $O/r2_${V}_bench.json 2> $O/r2_${V}_bench.err; echo "bench rc=$?"; grep -v "LOG(INFO)" $O/r2_${V}_bench.err | tail -6
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/r2_${V}_bench_reference.json 2>/dev/null; echo "ref rc=$?"
python - $O/r2_${V}_bench.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("config2", round(d["value"]/1e6,1), "M/s", round(d["ms_per_step"],3), "ms; e2e", round(d["e2e"]["ms_per_step"],2), d["stage_ms_per_step"], d["parity"]["sentences_mismatching"], "frac", round(d["roofline"]["frac"],4), "traffic", d["roofline"]["traffic"], "launches", d["gpu_launches"])
for k,v in d.get("configs",{}).items():
    if "error" in v: print(k, v); continue
    if k=="pieces":
        print(k, round(v["value"]/1e6,1), "M/s", round(v["ms_per_step"],3), "ms", v.get("stage_ms_per_step"), (v.get("parity") or {}).get("sentences_mismatching"), "wall", v.get("wall_s")); continue
    if k=="estep":
        for n in ("final","seed"): print(k,n, round(v[n]["value"]/1e6,1),"Mwords/s", round(v[n]["ms_per_step"],3), v[n]["parity"]["ok"])
    else: print(k, round(v["value"]/1e6,1), "M/s", round(v["ms_per_step"],3), "ms; e2e", round(v["e2e"]["ms_per_step"],2), v["parity"]["sentences_mismatching"], v["general_tier_sentences_per_step"], v["viterbi_tier_sentences_per_step"], "wall", v.get("wall_s"))
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_${V}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-configs > $O/r2_${V}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_memo_scan_words|k_memo_emit|k_compact' -c 3 -f -o $O/r2_${V}_full_config2 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-configs > $O/r2_${V}_ncu_full_config2.log 2>&1; echo "ncu full rc=$?"
