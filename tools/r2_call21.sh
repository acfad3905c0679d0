set -x
mkdir -p gpurun_out
B="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu --no-configs --workload P8 --samples 1024"
timeout 300 python bench.py $B > gpurun_out/r2_c21_P8_release.log 2>&1
cp concrete_b200/libconcrete_b200.so /tmp/orig.so
cp concrete_b200/libconcrete_b200_relaxed.so concrete_b200/libconcrete_b200.so
timeout 300 python bench.py $B > gpurun_out/r2_c21_P8_relaxed.log 2>&1
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "deterministic or cluster_rounds" > gpurun_out/r2_c21_pytest_relaxed.log 2>&1; tail -2 gpurun_out/r2_c21_pytest_relaxed.log
cp /tmp/orig.so concrete_b200/libconcrete_b200.so
grep -H -o '"launch_ms": [0-9.]*' gpurun_out/r2_c21_*.log
