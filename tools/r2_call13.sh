set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_wop.py -m gpu -x -q -k "two_streams or runtime_wrapper or deterministic" > gpurun_out/r2_pytest_c13.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_c13.log
tail -5 gpurun_out/r2_pytest_c13.log
B="--steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu --no-configs"
timeout 300 python bench.py $B > gpurun_out/r2_c13_s1.log 2>&1
for skew in 0 4000 8000 12000; do
  timeout 300 python bench.py $B --w32-streams 2 --w32-skew-ns $skew > gpurun_out/r2_c13_s2_$skew.log 2>&1
done
grep -H -o '"launch_ms": [0-9.]*' gpurun_out/r2_c13_s*.log
