set -x
# group kernel on the v4 key (k = 2, N = 1024, small batches): key rows prefetched into L1 at the start of the CMUX (pf = new, nopf = before)
mkdir -p gpurun_out
B="--steps 5 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
line() { echo "$1 $(grep -o '"launch_ms": [0-9.]*' $2 | head -1) $(grep -o '"verified_decryptions": [a-z]*' $2)"; }
cp concrete_b200/libconcrete_b200.so /tmp/lib_main.so
for v in pf nopf pf nopf; do
  if [ $v = nopf ]; then cp build/ab/libconcrete_b200_nopf.so concrete_b200/libconcrete_b200.so; else cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so; fi
  for S in 64 148 296; do
    timeout 300 python bench.py --workload P4 --samples $S $B > gpurun_out/r2_c47_${v}_P4_$S.log 2>&1
    line "$v P4 $S" gpurun_out/r2_c47_${v}_P4_$S.log
  done
done
cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "k2 or P4 or parameter_table or multi_level" > gpurun_out/r2_c47_pytest.log 2>&1; tail -2 gpurun_out/r2_c47_pytest.log
