set -x
for W in P7 P8 P6 P4; do
  A="--workload $W --samples 148 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-ref-gpu"
  timeout 300 python bench.py $A > gpurun_out/r2_plain_$W.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:pbs_kernel -s 3 -c 1 -o gpurun_out/r2_base_$W python bench.py $A > gpurun_out/r2_ncu_$W.log 2>&1
  bash tools/ncu_export.sh gpurun_out/r2_base_$W
done
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest0.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest0.log
du -sh gpurun_out
