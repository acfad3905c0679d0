set -x
# last single-stage pass reached by a lane-pair shuffle instead of a shared-memory exchange (shx1 = new default, shx0 = before):
# parity + determinism subset, then A/B at P8 and P4
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2_c34_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_c34_pytest.log
tail -3 gpurun_out/r2_c34_pytest.log
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
line() { echo "$1 $(grep -o '"launch_ms": [0-9.]*' $2 | head -1) $(grep -o '"verified_decryptions": [a-z]*' $2)"; }
cp concrete_b200/libconcrete_b200.so /tmp/lib_main.so
for v in shx1 shx0; do
  if [ $v = shx0 ]; then cp build/ab/libconcrete_b200_shx0.so concrete_b200/libconcrete_b200.so; else cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so; fi
  for W in "P8 1024" "P4 16384" "P4 740"; do set -- $W
    timeout 300 python bench.py --workload $1 --samples $2 $B > gpurun_out/r2_c34_${v}_$1_$2.log 2>&1
    line "$v $1 $2" gpurun_out/r2_c34_${v}_$1_$2.log
  done
done
cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so
