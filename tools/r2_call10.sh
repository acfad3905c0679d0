set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_seeded.py -m gpu -x -q -k "v4 or large_poly or deterministic or family or runtime_wrappers or pbs_parity or external" > gpurun_out/r2_pytest_c10.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_c10.log
B="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu"
for alt in 0 125 150 200; do
  timeout 600 python bench.py $B --workload P7 --samples 4096 --v4-alt $alt > gpurun_out/r2_c10_P7_alt$alt.log 2>&1
done
for alt in 0 125 150; do
  timeout 600 python bench.py $B --workload P6 --samples 4096 --v4-alt $alt > gpurun_out/r2_c10_P6_alt$alt.log 2>&1
done
timeout 600 python bench.py $B --workload P8 --samples 1024 --v4-alt 200 > gpurun_out/r2_c10_P8_alt200.log 2>&1
grep -H -o '"launch_ms": [0-9.]*' gpurun_out/r2_c10_*.log
tail -3 gpurun_out/r2_pytest_c10.log
