# One GPU pass of round 2: all GPU tests, the driver-shaped bench line, launch list.
# Usage (build container): gpurun --timeout 2700 -- 'bash tools/gpu_r2.sh v1'
V=${1:-vX}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader
nproc
( time timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -p no:cacheprovider ) > $O/r2_${V}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2_${V}_pytest_gpu.log
grep -v "LOG(INFO)" $O/r2_${V}_pytest_gpu.log | tail -40
( time timeout 900 python bench.py ) > $O/r2_${V}_bench.json 2> $O/r2_${V}_bench.err; echo "bench rc=$?"
tail -c 6000 $O/r2_${V}_bench.json; grep -v "LOG(INFO)" $O/r2_${V}_bench.err | tail -15
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/r2_${V}_bench_reference.json 2>/dev/null; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_${V}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-configs > $O/r2_${V}_ncu_launches.log 2>&1; echo "ncu rc=$?"
ls -la $O | tail -12
if [ -n "$2" ]; then
# full captures: the top kernels of config 2 and of config 4 (one launch each)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_memo_scan_words|k_memo_emit|k_compact' -c 3 -f -o $O/r2_${V}_full_config2 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-configs > $O/r2_${V}_ncu_full_config2.log 2>&1; echo "ncu full config2 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_encode_fast|k_encode_warp|k_normalize_fast|k_normalize_long' -c 4 -f -o $O/r2_${V}_full_config4 python bench.py --model ref_ja_8k --corpus wagahai --sentences 400000 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-configs > $O/r2_${V}_ncu_full_config4.log 2>&1; echo "ncu full config4 rc=$?"
ls -la $O | tail -6
fi
