# Round-2 pass B: all GPU tests, no-alloc test alone in a fresh process, full bench line, config 5 at N=1, ncu of emit/compact.
V=${1:-vX}
O=gpurun_out
mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -p no:cacheprovider ) > $O/r2_${V}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2_${V}_pytest_gpu.log
grep -v "LOG(INFO)" $O/r2_${V}_pytest_gpu.log | tail -30
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -2
( time timeout 900 python bench.py ) > $O/r2_${V}_bench.json 2> $O/r2_${V}_bench.err; echo "bench rc=$?"; grep -v "LOG(INFO)" $O/r2_${V}_bench.err | tail -8
python - $O/r2_${V}_bench.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("config2", round(d["value"]/1e6,1), "M/s", round(d["ms_per_step"],3), "ms; e2e", round(d["e2e"]["ms_per_step"],2), d["stage_ms_per_step"], d["parity"], d["roofline"]["frac"], d["roofline"]["traffic"])
for k,v in d.get("configs",{}).items():
    if "error" in v: print(k, v); continue
    if k=="estep":
        for n in ("final","seed"): print(k,n, round(v[n]["value"]/1e6,1),"Mwords/s", round(v[n]["ms_per_step"],3), v[n].get("parity"))
    else: print(k, round(v["value"]/1e6,1), "M/s", round(v["ms_per_step"],3), "ms; e2e", round(v["e2e"]["ms_per_step"],2), v["stage_ms_per_step"], v["parity"]["sentences_mismatching"], v["general_tier_sentences_per_step"], v["viterbi_tier_sentences_per_step"])
PY
( time timeout 600 python tools/config5.py --blocks 100 ) > $O/r2_${V}_config5_1gpu.json 2> $O/r2_${V}_config5_1gpu.err; echo "config5 rc=$?"; cat $O/r2_${V}_config5_1gpu.json; grep -v "LOG(INFO)" $O/r2_${V}_config5_1gpu.err | tail -5
timeout 600 python tools/config5.py --blocks 100 --threads 1 --no-pcie > $O/r2_${V}_config5_1gpu_serial.json 2>/dev/null; cat $O/r2_${V}_config5_1gpu_serial.json | cut -c1-400
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_memo_scan_words|k_memo_emit|k_compact' -c 3 -f -o $O/r2_${V}_full_config2 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-configs > $O/r2_${V}_ncu_full_config2.log 2>&1; echo "ncu rc=$?"
