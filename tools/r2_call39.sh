set -x
# 32 points per thread (8 warps, 255 registers, twiddles in tensor memory) for N = 4096 / 8192: parity, then A/B against the defaults
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tuning_variants or large_poly or external_product or fourier" > gpurun_out/r2_c39_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_c39_pytest.log
tail -3 gpurun_out/r2_c39_pytest.log
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
line() { echo "$1 $(grep -o '"launch_ms": [0-9.]*' $2 | head -1) $(grep -o '"verified_decryptions": [a-z]*' $2)"; }
for v in "alt0" "alt4 --v4-alt 4"; do set -- $v; tag=$1; shift
  for W in "P6 4096" "P7 4096" "P6 296" "P7 148"; do set -- $W "$@"; w=$1; s=$2; shift 2
    timeout 300 python bench.py --workload $w --samples $s $B "$@" > gpurun_out/r2_c39_${tag}_${w}_$s.log 2>&1
    line "$tag $w $s" gpurun_out/r2_c39_${tag}_${w}_$s.log
  done
done
