set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_last_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_last_pytest.log
tail -3 gpurun_out/r2_last_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 600 python bench.py --steps 3 --warmup 3 --no-configs > gpurun_out/r2_last_bench.log 2>&1; tail -c 300 gpurun_out/r2_last_bench.log
