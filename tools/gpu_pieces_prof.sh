V=${1:-vX}; O=gpurun_out; mkdir -p $O
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/r2_${V}_launches_pieces.csv python tools/pieces_bench.py --op pieces --sentences 10000000 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $O/r2_${V}_ncu_pieces.log 2>&1; echo "ncu rc=$?"
