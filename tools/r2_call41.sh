set -x
mkdir -p gpurun_out
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
line() { echo "$1 $(grep -o '"launch_ms": [0-9.]*' $2 | head -1) $(grep -o '"verified_decryptions": [a-z]*' $2)"; }
for v in "alt0" "alt5 --v4-alt 5"; do set -- $v; tag=$1; shift
  for W in "P6 4096" "P7 4096"; do set -- $W "$@"; w=$1; s=$2; shift 2
    timeout 300 python bench.py --workload $w --samples $s $B "$@" > gpurun_out/r2_c41_${tag}_${w}_$s.log 2>&1
    line "$tag $w $s" gpurun_out/r2_c41_${tag}_${w}_$s.log
  done
done
