set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "v4 or family or large_poly or pbs_parity" > gpurun_out/r2_pytest_v4c.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_v4c.log
timeout 600 python -m pytest tests/test_gpu_sdfg.py -m gpu -q > gpurun_out/r2_pytest_sdfg.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_sdfg.log
B="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu"
for W in "P7 4096" "P6 4096" "P4 16384" "P4 64"; do set -- $W
  timeout 300 python bench.py --workload $1 --samples $2 $B > gpurun_out/r2_v4c_$1_$2.log 2>&1
done
timeout 300 python bench.py --workload P2048 --samples 16384 --pbs-plan 4 $B > gpurun_out/r2_v4c_P2048_plan4.log 2>&1
timeout 900 python -m pytest tests/test_gpu_workloads.py -m gpu -x -q -k "noise_variance or cfg3" > gpurun_out/r2_pytest_wl.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_wl.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/r2_smoke.log
