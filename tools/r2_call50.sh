set -x
# P4 on v4: the whole GGSW as ONE ring block per CMUX (--v4-alt 6, plan 4) against the default two half-GGSW blocks
mkdir -p gpurun_out
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
line() { echo "$1 $(grep -o '"launch_ms": [0-9.]*' $2 | head -1) $(grep -o '"verified_decryptions": [a-z]*' $2)"; }
for v in "alt0 --pbs-plan 4" "alt6 --pbs-plan 4 --v4-alt 6" "alt0b --pbs-plan 4" "alt6b --pbs-plan 4 --v4-alt 6"; do set -- $v; tag=$1; shift
  for S in 16384 740; do
    timeout 300 python bench.py --workload P4 --samples $S $B "$@" > gpurun_out/r2_c50_${tag}_P4_$S.log 2>&1
    line "$tag P4 $S" gpurun_out/r2_c50_${tag}_P4_$S.log
  done
done
