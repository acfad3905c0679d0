set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "v4 or family or large_poly or pbs_parity" > gpurun_out/r2_pytest_v4b.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_v4b.log
B="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu"
for W in "P7 4096" "P6 4096" "P4 16384" "P4 64"; do set -- $W
  timeout 300 python bench.py --workload $1 --samples $2 $B > gpurun_out/r2_v4b_$1_$2.log 2>&1
done
timeout 300 python bench.py --workload P2048 --samples 16384 --pbs-plan 4 $B > gpurun_out/r2_v4b_P2048_plan4.log 2>&1
