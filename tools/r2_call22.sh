set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_seeded.py -m gpu -x -q -k "deterministic or cluster or v4 or large_poly or P8" > gpurun_out/r2_c22_pytest.log 2>&1; tail -2 gpurun_out/r2_c22_pytest.log
B="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu --no-configs --workload P8"
timeout 300 python bench.py $B --samples 1024 > gpurun_out/r2_c22_P8_1024.log 2>&1
timeout 300 python bench.py $B --samples 4096 > gpurun_out/r2_c22_P8_4096.log 2>&1
grep -H -o '"launch_ms": [0-9.]*' gpurun_out/r2_c22_*.log
