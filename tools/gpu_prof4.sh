# parity subset, config 4 bench at 20 k / 200 k / 2.2 M lines, then ncu --set full of k_encode_warp at 20 k lines.
V=${1:-vX}; O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -p no:cacheprovider -k "corpus_parity or long_lines or near_tie or encode_fast_kernel or edge_sentences or option_paths or host_chunk or test_smoke" ) > $O/r2_${V}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2_${V}_pytest_gpu.log
grep -v "LOG(INFO)" $O/r2_${V}_pytest_gpu.log | tail -5
for n in 20000 200000 2200000; do
timeout 600 python bench.py --cpu-sample 60000 --no-configs --model ref_ja_8k --corpus wagahai --sentences $n > $O/r2_${V}_bench_c4_$n.json 2> $O/r2_${V}_bench_c4_$n.err; echo "bench $n rc=$?"
python - $O/r2_${V}_bench_c4_$n.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(round(d["value"]/1e6,2), "M lines/s", round(d["ms_per_step"],3), "ms", round(d["input_gb_per_s"],2),"GB/s e2e", round(d.get("e2e",{}).get("ms_per_step",0),2), d.get("stage_ms_per_step"), (d.get("parity") or {}).get("sentences_mismatching"), d["viterbi_tier_sentences_per_step"], d["general_tier_sentences_per_step"])
except Exception as e: print("bad", e)
PY
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_${V}_launches_c4_2200000.csv python bench.py --model ref_ja_8k --corpus wagahai --sentences 2200000 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-configs > $O/r2_${V}_ncu_l.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_encode_warp' -c 1 -f -o $O/r2_${V}_full_warp_20k python bench.py --model ref_ja_8k --corpus wagahai --sentences 20000 --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-configs > $O/r2_${V}_ncu_warp.log 2>&1; echo "ncu rc=$?"
