# One GPU pass of the round: tests, bench lines of every config, launch list, ncu --set full captures.
# Usage (build container): gpurun --timeout 2400 -- 'bash tools/gpu_round.sh v11'
V=${1:-vX}
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/${V}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${V}_pytest_gpu.log; tail -2 $O/${V}_pytest_gpu.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${V}_bench_reference.json 2> $O/${V}_bench_reference.err
timeout 600 python bench.py > $O/${V}_bench_config2.json 2> $O/${V}_bench_config2.err; echo "bench rc=$?"
timeout 600 python bench.py --corpus syn_dirty --no-cpu-baseline > $O/${V}_bench_config2_dirty.json 2>/dev/null
timeout 600 python bench.py --model syn_bpe32k > $O/${V}_bench_config3.json 2>/dev/null
timeout 600 python bench.py --model ref_ja_8k --corpus wagahai --sentences 200000 --steps 3 > $O/${V}_bench_config4.json 2>/dev/null
timeout 600 python tools/pieces_bench.py --op pieces --sentences 10000000 > $O/${V}_pieces_config2.json 2>/dev/null
timeout 600 python tools/pieces_bench.py --op pieces --sentences 10000000 --model syn_bpe32k > $O/${V}_pieces_config3.json 2>/dev/null
timeout 600 python tools/pieces_bench.py --op pieces --model ref_ja_8k --corpus wagahai --sentences 200000 > $O/${V}_pieces_config4.json 2>/dev/null
timeout 600 python tools/pieces_bench.py --op decode --sentences 10000000 > $O/${V}_decode_config2.json 2>/dev/null
timeout 600 python tools/pieces_bench.py --op dense --sentences 10000000 > $O/${V}_dense_config2.json 2>/dev/null
# launch lists (per-launch times are cold-cache and serialised: shares, not absolutes)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${V}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > $O/${V}_ncu_launches.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${V}_launches_pieces.csv python tools/pieces_bench.py --op pieces --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $O/${V}_ncu_launches_pieces.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/${V}_launches_decode.csv python tools/pieces_bench.py --op decode --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $O/${V}_ncu_launches_decode.log 2>&1
# full captures: the top kernels of the id path, the pieces path and decode (one launch each)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_memo_scan_words|k_memo_emit|k_compact|k_encode_warp' -c 4 -f -o $O/${V}_full_ids python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $O/${V}_ncu_full_ids.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_pc_normalize|k_pc_spans|k_pc_copy_norm' -c 3 -f -o $O/${V}_full_pieces python tools/pieces_bench.py --op pieces --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $O/${V}_ncu_full_pieces.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_decode' -c 2 -f -o $O/${V}_full_decode python tools/pieces_bench.py --op decode --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $O/${V}_ncu_full_decode.log 2>&1
ls -la $O | tail -40
