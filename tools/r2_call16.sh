set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_c16.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_c16.log
tail -4 gpurun_out/r2_pytest_c16.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_1gpu.log 2>&1; echo "bench rc=$?"
A="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu --no-configs"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_launches.csv python bench.py $A > gpurun_out/r2_ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ks_umma_kernel -s 2 -c 1 -o gpurun_out/r2_ks_umma python bench.py $A > gpurun_out/r2_ncu_ks.log 2>&1
bash tools/ncu_export.sh gpurun_out/r2_ks_umma
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pbs_kernel_w32v3 -s 3 -c 1 -o gpurun_out/r2_w32v3_16k python bench.py $A > gpurun_out/r2_ncu_v3_16k.log 2>&1
bash tools/ncu_export.sh gpurun_out/r2_w32v3_16k
du -sh gpurun_out
