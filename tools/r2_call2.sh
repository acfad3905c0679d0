set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "v4 or family or more_than or cluster or pbs_parity or external_product or large_poly" > gpurun_out/r2_pytest_v4.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_v4.log
B="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu"
for W in "P7 4096" "P6 4096" "P4 16384" "P4 64"; do set -- $W
  timeout 300 python bench.py --workload $1 --samples $2 $B > gpurun_out/r2_v4_$1_$2.log 2>&1
done
timeout 300 python bench.py --workload P2048 --samples 16384 --pbs-plan 4 $B > gpurun_out/r2_v4_P2048_plan4.log 2>&1
timeout 300 python bench.py --workload P2048 --samples 16384 $B > gpurun_out/r2_v3_P2048.log 2>&1
