set -x
# no-alloc keyswitch entry point: alone in a fresh process, then after the whole GPU suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2_no_alloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/r2_no_alloc_alone.log
tail -5 gpurun_out/r2_no_alloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_with_no_alloc.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pytest_with_no_alloc.log
tail -5 gpurun_out/r2_pytest_with_no_alloc.log
