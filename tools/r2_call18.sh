set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "family or v4" > gpurun_out/r2_pytest_c18.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_c18.log
tail -4 gpurun_out/r2_pytest_c18.log
B="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu --no-configs --workload P4"
timeout 300 python bench.py $B --samples 16384 --pbs-plan 4 > gpurun_out/r2_c18_P4_v4p8_16384.log 2>&1
timeout 300 python bench.py $B --samples 64 --pbs-plan 4 > gpurun_out/r2_c18_P4_v4p8_64.log 2>&1
timeout 300 python bench.py $B --samples 4096 --pbs-plan 4 > gpurun_out/r2_c18_P4_v4p8_4096.log 2>&1
grep -H -o '"launch_ms": [0-9.]*' gpurun_out/r2_c18_*.log
