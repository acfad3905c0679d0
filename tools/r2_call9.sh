set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_seeded.py -m gpu -x -q > gpurun_out/r2_pytest_seeded.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_seeded.log
tail -20 gpurun_out/r2_pytest_seeded.log
