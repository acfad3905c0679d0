set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "v4 or large_poly or cluster or deterministic or family" > gpurun_out/r2_pytest_p8b.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_p8b.log
B="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu --workload P8 --samples 1024"
for alt in 0 1 2 3; do
  timeout 600 python bench.py $B --v4-alt $alt > gpurun_out/r2_bench_P8_alt$alt.log 2>&1
done
timeout 300 python bench.py --workload P4 --samples 16384 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu > gpurun_out/r2_bench_P4_grp.log 2>&1
grep -h -o '"launch_ms": [0-9.]*' gpurun_out/r2_bench_P8_alt*.log gpurun_out/r2_bench_P4_grp.log
