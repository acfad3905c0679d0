set -x
# L2 evict_last hint on key fills (keep1 = new default, keep0 = without, v3 only differs): DRAM bytes per launch and time at the bench configuration
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "parity or family or determin or k2" > gpurun_out/r2_c32_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_c32_pytest.log
tail -3 gpurun_out/r2_c32_pytest.log
B="--steps 5 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
A="--steps 1 --warmup 1 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
cp concrete_b200/libconcrete_b200.so /tmp/lib_main.so
for v in keep1 keep0; do
  if [ $v = keep0 ]; then cp build/ab/libconcrete_b200_keep0.so concrete_b200/libconcrete_b200.so; else cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so; fi
  timeout 300 python bench.py $B > gpurun_out/r2_c32_${v}_bench.log 2>&1
  echo "$v $(grep -o '"launch_ms": [0-9.]*' gpurun_out/r2_c32_${v}_bench.log | head -1) $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_c32_${v}_bench.log) $(grep -o '"verified_decryptions": [a-z]*' gpurun_out/r2_c32_${v}_bench.log)"
  timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:pbs_kernel_w32v3 -s 1 -c 1 --csv --log-file gpurun_out/r2_c32_${v}_dram.csv python bench.py $A > gpurun_out/r2_c32_${v}_ncu.log 2>&1
  grep -E "dram__bytes|gpu__time|lts__" gpurun_out/r2_c32_${v}_dram.csv | cut -d, -f5,13-
done
cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so
# P4 with the hint in the v4 kernel (59 MB key): time only
timeout 300 python bench.py --workload P4 --samples 16384 $B > gpurun_out/r2_c32_P4.log 2>&1
echo "P4 16384 $(grep -o '"launch_ms": [0-9.]*' gpurun_out/r2_c32_P4.log | head -1)"
