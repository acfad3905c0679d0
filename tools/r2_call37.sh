set -x
mkdir -p gpurun_out
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
line() { echo "$1 $(grep -o '"launch_ms": [0-9.]*' $2 | head -1) $(grep -o '"verified_decryptions": [a-z]*' $2)"; }
for v in "alt0" "alt2 --v4-alt 2"; do set -- $v; tag=$1; shift
  for S in 1024 74; do
    timeout 300 python bench.py --workload P8 --samples $S $B "$@" > gpurun_out/r2_c37_${tag}_P8_$S.log 2>&1
    line "$tag P8 $S" gpurun_out/r2_c37_${tag}_P8_$S.log
  done
done
