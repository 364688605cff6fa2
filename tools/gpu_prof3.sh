# launch list of config 4 at 20 k and 200 k lines + ncu --set full of k_encode_warp at 20 k lines (tail only).
V=${1:-vX}; O=gpurun_out; mkdir -p $O
for n in 20000 200000; do
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_${V}_launches_c4_$n.csv python bench.py --model ref_ja_8k --corpus wagahai --sentences $n --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-configs > $O/r2_${V}_ncu_l_$n.log 2>&1; echo "ncu launches $n rc=$?"
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_encode_warp' -c 1 -f -o $O/r2_${V}_full_warp_20k python bench.py --model ref_ja_8k --corpus wagahai --sentences 20000 --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-configs > $O/r2_${V}_ncu_warp.log 2>&1; echo "ncu rc=$?"
