set -x
# fresh --set full capture of the bench kernel (final: 6-op inverse butterflies) at the bench configuration
mkdir -p gpurun_out
A="--steps 1 --warmup 1 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pbs_kernel_w32v3 -s 1 -c 1 -o gpurun_out/r2_final_w32v3_16k python bench.py $A > gpurun_out/r2_final_ncu_w32v3.log 2>&1
bash tools/ncu_export.sh gpurun_out/r2_final_w32v3_16k
rm -f gpurun_out/*.source.csv.gz
ls -la gpurun_out | grep w32v3
