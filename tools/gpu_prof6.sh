# ncu --set full of the pieces-path kernels (one chunk of 1 M config-2 sentences)
V=${1:-vX}; O=gpurun_out; mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_pc_' -c 5 -f -o $O/r2_${V}_full_pieces python tools/pieces_bench.py --op pieces --sentences 1000000 --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > $O/r2_${V}_ncu_pieces.log 2>&1; echo "ncu rc=$?"
