set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_wop.py -m gpu -x -q > gpurun_out/r2_pytest_wop.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_wop.log
tail -40 gpurun_out/r2_pytest_wop.log
