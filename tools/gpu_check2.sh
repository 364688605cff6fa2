set -x
timeout 900 python -m pytest tests/test_pieces.py -m gpu -x -q > gpurun_out/v9_pytest_pieces.log 2>&1; echo "pytest rc=$?" >> gpurun_out/v9_pytest_pieces.log
tail -30 gpurun_out/v9_pytest_pieces.log
