set -x
# bench kernel (N = 2048): two compute groups per key fill (gpf2 = new, gpf1 = before): parity on the new library, A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "P2048 or parity or determin or zero_mask or many_lut or golden or chain or legacy or amortized or memref or full_size or external" > gpurun_out/r2_c52_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_c52_pytest.log
tail -3 gpurun_out/r2_c52_pytest.log
B="--steps 5 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
line() { echo "$1 $(grep -o '"launch_ms": [0-9.]*' $2 | head -1) $(grep -o '"verified_decryptions": [a-z]*' $2)"; }
cp concrete_b200/libconcrete_b200.so /tmp/lib_main.so
for v in gpf2 gpf1 gpf2 gpf1; do
  if [ $v = gpf1 ]; then cp build/ab/libconcrete_b200_gpf1.so concrete_b200/libconcrete_b200.so; else cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so; fi
  timeout 300 python bench.py $B > gpurun_out/r2_c52_${v}.log 2>&1
  line "$v P2048 16384" gpurun_out/r2_c52_${v}.log
  timeout 300 python bench.py --samples 888 $B > gpurun_out/r2_c52_${v}_888.log 2>&1
  line "$v P2048 888" gpurun_out/r2_c52_${v}_888.log
done
cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so
