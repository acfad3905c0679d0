set -x
# product phase (with the unroll pragma back in place): main = early count/probe + batched loads, e0 = batched loads only, base = neither
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2_c44_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_c44_pytest.log
tail -3 gpurun_out/r2_c44_pytest.log
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
line() { echo "$1 $(grep -o '"launch_ms": [0-9.]*' $2 | head -1) $(grep -o '"verified_decryptions": [a-z]*' $2)"; }
cp concrete_b200/libconcrete_b200.so /tmp/lib_main.so
for v in main e0 base; do
  if [ $v = main ]; then cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so; else cp build/ab/libconcrete_b200_$v.so concrete_b200/libconcrete_b200.so; fi
  for W in "P4 16384" "P6 4096" "P7 4096" "P8 1024"; do set -- $W
    timeout 300 python bench.py --workload $1 --samples $2 $B > gpurun_out/r2_c44_${v}_$1_$2.log 2>&1
    line "$v $1 $2" gpurun_out/r2_c44_${v}_$1_$2.log
  done
done
cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so
