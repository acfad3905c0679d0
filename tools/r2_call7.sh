set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "v4 or large_poly or cluster or external_product" > gpurun_out/r2_pytest_p8.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_p8.log
B="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu"
timeout 600 python bench.py --workload P8 --samples 1024 $B > gpurun_out/r2_bench_v4cl_P8_1024.log 2>&1
timeout 600 python bench.py --workload P8 --samples 4096 $B > gpurun_out/r2_bench_v4cl_P8_4096.log 2>&1
A="--workload P8 --samples 74 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-ref-gpu"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pbs_kernel -s 3 -c 1 -o gpurun_out/r2_v4cl_P8 python bench.py $A > gpurun_out/r2_ncu_v4cl_P8.log 2>&1
bash tools/ncu_export.sh gpurun_out/r2_v4cl_P8
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/r2_bench_configs.log 2>&1
