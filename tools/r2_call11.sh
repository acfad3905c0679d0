set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_c11.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_c11.log
B="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu"
timeout 600 python bench.py $B --workload P4 --samples 16384 > gpurun_out/r2_c11_P4_16384.log 2>&1
timeout 600 python bench.py $B --workload P4 --samples 64 > gpurun_out/r2_c11_P4_64.log 2>&1
grep -H -o '"launch_ms": [0-9.]*' gpurun_out/r2_c11_*.log
tail -3 gpurun_out/r2_pytest_c11.log
