set -x
# k = 2, N = 1024 (P4): v4 warp-per-polynomial kernel for batches >= 740, group kernel on the same key below.
# Full GPU suite, then P4 at 64 / 739 / 740 / 4096 / 16384 on the automatic plan, ncu of the new P4 kernel.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c30_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_c30_pytest.log
tail -4 gpurun_out/r2_c30_pytest.log
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e"
for S in 64 739 740 4096 16384; do
  timeout 300 python bench.py --workload P4 --samples $S $B > gpurun_out/r2_c30_P4_$S.log 2>&1
  echo "P4 $S $(grep -o '"launch_ms": [0-9.]*' gpurun_out/r2_c30_P4_$S.log | head -1) $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_c30_P4_$S.log) $(grep -o '"verified_decryptions": [a-z]*' gpurun_out/r2_c30_P4_$S.log) $(grep -o '"vs_reference_gpu": [0-9.a-z]*' gpurun_out/r2_c30_P4_$S.log)"
done
A="--steps 1 --warmup 1 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pbs_kernel -s 3 -c 1 -o gpurun_out/r2_c30_P4 python bench.py --workload P4 --samples 740 $A > gpurun_out/r2_c30_ncu_P4.log 2>&1
bash tools/ncu_export.sh gpurun_out/r2_c30_P4
rm -f gpurun_out/*.source.csv.gz
