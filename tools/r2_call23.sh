set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_final.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_final.log
tail -3 gpurun_out/r2_pytest_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke_final.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r2_smoke_final.log; tail -2 gpurun_out/r2_smoke_final.log
timeout 900 python bench.py --steps 10 --warmup 5 > gpurun_out/r2_bench_final.log 2>&1; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r2_bench_final_ref.log 2>&1
B="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-configs"
for W in "P4 16384" "P4 4096" "P4 64" "P6 4096" "P7 4096" "P8 1024" "P8 4096"; do set -- $W
  timeout 900 python bench.py --workload $1 --samples $2 $B > gpurun_out/r2_final_$1_$2.log 2>&1
done
grep -H -o '"launch_ms": [0-9.]*' gpurun_out/r2_final_*.log
