set -x
# v4: twiddles of the non-first passes in tensor memory: parity suite, then A/B (tw1 = new, tw0 = global loads)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_tw_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_tw_pytest.log
tail -4 gpurun_out/r2_tw_pytest.log
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
cp concrete_b200/libconcrete_b200.so /tmp/lib_new.so
for v in tw1 tw0; do
  if [ $v = tw0 ]; then cp build/ab/libconcrete_b200_tw0.so concrete_b200/libconcrete_b200.so; else cp /tmp/lib_new.so concrete_b200/libconcrete_b200.so; fi
  for W in "P7 4096" "P6 4096" "P7 148" "P6 296"; do set -- $W
    timeout 300 python bench.py --workload $1 --samples $2 $B > gpurun_out/r2_tw_bench_${v}_$1_$2.log 2>&1
    echo "$v $1 $2 $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_tw_bench_${v}_$1_$2.log)"
  done
  timeout 300 python bench.py --workload P2048 --samples 16384 --pbs-plan 4 $B > gpurun_out/r2_tw_bench_${v}_P2048v4.log 2>&1
  echo "$v P2048 plan4 $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_tw_bench_${v}_P2048v4.log)"
done
cp /tmp/lib_new.so concrete_b200/libconcrete_b200.so
