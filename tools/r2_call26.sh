set -x
# v3 inverse pass 1 with 6-op butterflies: parity suite on the new library, then A/B bench (lf1 = new, lf0 = 8-op)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_lf_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_lf_pytest.log
tail -4 gpurun_out/r2_lf_pytest.log
B="--steps 5 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
cp concrete_b200/libconcrete_b200.so /tmp/lib_lf1.so
for rep in 1 2; do
  for v in lf1 lf0; do
    if [ $v = lf0 ]; then cp build/ab/libconcrete_b200_lf0.so concrete_b200/libconcrete_b200.so; else cp /tmp/lib_lf1.so concrete_b200/libconcrete_b200.so; fi
    timeout 300 python bench.py $B > gpurun_out/r2_lf_bench_${v}_$rep.log 2>&1
    grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_lf_bench_${v}_$rep.log
  done
done
cp /tmp/lib_lf1.so concrete_b200/libconcrete_b200.so
