set -x
# round-2 state of the tree on a fresh box: whole GPU suite, smoke, every BASELINE config, ncu of the v4/v3 kernels
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_full.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_full.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/r2_smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_P2048.log 2>&1
B="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 python bench.py --workload P2048 --samples 16384 --pbs-plan 4 $B --no-ref-gpu > gpurun_out/r2_bench_P2048_plan4.log 2>&1
for W in "P4 16384" "P4 4096" "P4 64" "P6 4096" "P7 4096" "P8 1024"; do set -- $W
  timeout 600 python bench.py --workload $1 --samples $2 $B > gpurun_out/r2_bench_$1_$2.log 2>&1
done
for W in "P2048 888 w32v3" "P7 148 v4" "P6 296 v4" "P4 740 v4" "P8 74 cl"; do set -- $W
  A="--workload $1 --samples $2 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-ref-gpu"
  timeout 300 python bench.py $A > gpurun_out/r2_plain_$1.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:pbs_kernel -s 3 -c 1 -o gpurun_out/r2_$3_$1 python bench.py $A > gpurun_out/r2_ncu_$1.log 2>&1
  bash tools/ncu_export.sh gpurun_out/r2_$3_$1
done
du -sh gpurun_out
