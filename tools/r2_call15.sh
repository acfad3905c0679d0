set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tensor_core_kernels" > gpurun_out/r2_pytest_c15.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_c15.log
tail -15 gpurun_out/r2_pytest_c15.log
B="--steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu --no-configs"
timeout 300 python bench.py $B --ks-plan 3 > gpurun_out/r2_c15_umma.log 2>&1
timeout 300 python bench.py $B --workload P8 --samples 1024 --ks-plan 3 > gpurun_out/r2_c15_umma_P8.log 2>&1
grep -H -o '"launch_ms": [0-9.]*' gpurun_out/r2_c15_*.log
