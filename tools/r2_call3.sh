set -x
for W in "P7 148" "P4 740"; do set -- $W
  A="--workload $1 --samples $2 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-ref-gpu"
  timeout 300 python bench.py $A > gpurun_out/r2_plain_v4_$1.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:pbs_kernel -s 3 -c 1 -o gpurun_out/r2_v4a_$1 python bench.py $A > gpurun_out/r2_ncu_v4_$1.log 2>&1
  bash tools/ncu_export.sh gpurun_out/r2_v4a_$1
done
