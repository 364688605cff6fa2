# ncu --set full of k_encode_warp (config 4, 200 k lines) for two group sizes.
V=${1:-vX}; O=gpurun_out; mkdir -p $O
for g in 32 8; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_encode_warp' -c 1 -f -o $O/r2_${V}_full_warp_g$g python bench.py --model ref_ja_8k --corpus wagahai --sentences 200000 --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-configs --option warp_group=$g > $O/r2_${V}_ncu_warp_g$g.log 2>&1; echo "ncu g$g rc=$?"; tail -2 $O/r2_${V}_ncu_warp_g$g.log
done
