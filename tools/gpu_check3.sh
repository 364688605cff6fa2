set -x
timeout 900 python -m pytest tests/test_pieces.py -m gpu -x -q > gpurun_out/v9_pytest_pieces.log 2>&1; echo "pytest rc=$?" >> gpurun_out/v9_pytest_pieces.log
tail -3 gpurun_out/v9_pytest_pieces.log
for op in pieces decode dense; do
  timeout 600 python tools/pieces_bench.py --op $op > gpurun_out/v9_${op}_config2.json 2> gpurun_out/v9_${op}_config2.err; echo "$op rc=$?"; cat gpurun_out/v9_${op}_config2.json; tail -3 gpurun_out/v9_${op}_config2.err
done
timeout 600 python tools/pieces_bench.py --op pieces --model syn_bpe32k > gpurun_out/v9_pieces_config3.json 2> gpurun_out/v9_pieces_config3.err; cat gpurun_out/v9_pieces_config3.json
timeout 600 python tools/pieces_bench.py --op pieces --model ref_ja_8k --corpus wagahai --sentences 200000 > gpurun_out/v9_pieces_config4.json 2> gpurun_out/v9_pieces_config4.err; cat gpurun_out/v9_pieces_config4.json
