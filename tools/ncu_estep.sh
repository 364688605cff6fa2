O=gpurun_out; mkdir -p $O
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_estep' -c 2 -f -o $O/r2_v5_full_estep python tools/estep_bench.py --no-reference --steps 1 --warmup 1 > $O/r2_v5_ncu_estep.log 2>&1; echo "ncu rc=$?"; tail -3 $O/r2_v5_ncu_estep.log
