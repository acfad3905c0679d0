set -x
# v4 product phase pipelined (pipe1 = new default, pipe0 = first form): parity, A/B on P4/P6/P7; P4 kernel switch around 297
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2_c31_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_c31_pytest.log
tail -3 gpurun_out/r2_c31_pytest.log
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
line() { echo "$1 $(grep -o '"launch_ms": [0-9.]*' $2 | head -1) $(grep -o '"ms_per_step": [0-9.]*' $2) $(grep -o '"verified_decryptions": [a-z]*' $2)"; }
cp concrete_b200/libconcrete_b200.so /tmp/lib_main.so
for v in pipe1 pipe0; do
  if [ $v = pipe0 ]; then cp build/ab/libconcrete_b200_pipe0.so concrete_b200/libconcrete_b200.so; else cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so; fi
  for W in "P4 16384" "P6 4096" "P7 4096" "P4 740"; do set -- $W
    timeout 300 python bench.py --workload $1 --samples $2 $B > gpurun_out/r2_c31_${v}_$1_$2.log 2>&1
    line "$v $1 $2" gpurun_out/r2_c31_${v}_$1_$2.log
  done
done
cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so
for S in 64 148 296 297 444; do
  for P in 0 2 4; do
    timeout 300 python bench.py --workload P4 --samples $S --pbs-plan $P $B > gpurun_out/r2_c31_P4_${S}_plan$P.log 2>&1
    line "P4 $S plan$P" gpurun_out/r2_c31_P4_${S}_plan$P.log
  done
done
