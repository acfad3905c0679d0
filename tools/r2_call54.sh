set -x
# ring blocks that do not divide the rows of a level: 2 slots of 48 KiB (--v4-alt 7) against 3 slots of 32 KiB (default)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tuning_variants or large_poly or multi_level or cluster or determin" > gpurun_out/r2_c54_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_c54_pytest.log
tail -3 gpurun_out/r2_c54_pytest.log
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
line() { echo "$1 $(grep -o '"launch_ms": [0-9.]*' $2 | head -1) $(grep -o '"verified_decryptions": [a-z]*' $2)"; }
for v in "alt0" "alt7 --v4-alt 7"; do set -- $v; tag=$1; shift
  for W in "P6 4096" "P7 4096" "P8 1024"; do set -- $W "$@"; w=$1; s=$2; shift 2
    timeout 300 python bench.py --workload $w --samples $s $B "$@" > gpurun_out/r2_c54_${tag}_${w}_$s.log 2>&1
    line "$tag $w $s" gpurun_out/r2_c54_${tag}_${w}_$s.log
  done
done
