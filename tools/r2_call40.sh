set -x
# after removing the 32-point one-level variants: parity of test_gpu_parity.py; P8 default with / without twiddle prefetch (pf1 / main)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2_c40_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_c40_pytest.log
tail -3 gpurun_out/r2_c40_pytest.log
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
line() { echo "$1 $(grep -o '"launch_ms": [0-9.]*' $2 | head -1) $(grep -o '"verified_decryptions": [a-z]*' $2)"; }
cp concrete_b200/libconcrete_b200.so /tmp/lib_main.so
for v in main pf1; do
  if [ $v = pf1 ]; then cp build/ab/libconcrete_b200_pf1.so concrete_b200/libconcrete_b200.so; else cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so; fi
  timeout 300 python bench.py --workload P8 --samples 1024 $B > gpurun_out/r2_c40_${v}_P8.log 2>&1
  line "$v P8 1024" gpurun_out/r2_c40_${v}_P8.log
done
cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so
