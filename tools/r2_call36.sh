set -x
# two-level kernels: decomposition states parked in tensor memory instead of spilled (sttm1 = new default, sttm0 = before)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_workloads.py -m gpu -x -q > gpurun_out/r2_c36_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_c36_pytest.log
tail -3 gpurun_out/r2_c36_pytest.log
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
line() { echo "$1 $(grep -o '"launch_ms": [0-9.]*' $2 | head -1) $(grep -o '"verified_decryptions": [a-z]*' $2)"; }
cp concrete_b200/libconcrete_b200.so /tmp/lib_main.so
for v in sttm1 sttm0; do
  if [ $v = sttm0 ]; then cp build/ab/libconcrete_b200_sttm0.so concrete_b200/libconcrete_b200.so; else cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so; fi
  timeout 300 python bench.py --workload P8 --samples 1024 $B > gpurun_out/r2_c36_${v}_P8.log 2>&1
  line "$v P8 1024" gpurun_out/r2_c36_${v}_P8.log
done
cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so
A="--steps 1 --warmup 1 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pbs_kernel -s 3 -c 1 -o gpurun_out/r2_c36_P8 python bench.py --workload P8 --samples 74 $A > gpurun_out/r2_c36_ncu_P8.log 2>&1
bash tools/ncu_export.sh gpurun_out/r2_c36_P8
rm -f gpurun_out/*.source.csv.gz
