set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/v8_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/v8_pytest.log
timeout 600 python bench.py > gpurun_out/v8_bench_config2.json 2> gpurun_out/v8_bench_config2.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/v8_bench_reference.json 2> gpurun_out/v8_bench_reference.err
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/v8_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/v8_ncu.log 2>&1
tail -3 gpurun_out/v8_pytest.log; cat gpurun_out/v8_bench_config2.json
