set -x
mkdir -p gpurun_out
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
line() { echo "$1 $(grep -o '"launch_ms": [0-9.]*' $2 | head -1) $(grep -o '"verified_decryptions": [a-z]*' $2)"; }
cp concrete_b200/libconcrete_b200.so /tmp/lib_main.so
for v in main base main base; do
  if [ $v = base ]; then cp build/ab/libconcrete_b200_base.so concrete_b200/libconcrete_b200.so; else cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so; fi
  timeout 300 python bench.py --workload P8 --samples 1024 $B > gpurun_out/r2_c43_${v}_P8.log 2>&1
  line "$v P8 1024" gpurun_out/r2_c43_${v}_P8.log
done
cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so
timeout 300 python bench.py --workload P7 --samples 4096 $B > gpurun_out/r2_c43_main_P7.log 2>&1; line "main P7 4096" gpurun_out/r2_c43_main_P7.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_level or cluster or large_poly or tuning or k2" > gpurun_out/r2_c43_pytest.log 2>&1; tail -2 gpurun_out/r2_c43_pytest.log
