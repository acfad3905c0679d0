set -x
# DMMA probe; P4 (k = 2, N = 1024) on the group kernel vs v4 (8 points/thread, 3 ct/CTA) vs v4 alt 3 (warp per polynomial, 5 ct/CTA)
mkdir -p gpurun_out
./experiments/dmma_probe > gpurun_out/r2_c29_dmma_probe.json 2>&1; cat gpurun_out/r2_c29_dmma_probe.json
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
for S in 16384 64 4096; do
  for v in "grp" "v4 --pbs-plan 4" "v4alt3 --pbs-plan 4 --v4-alt 3"; do set -- $v; tag=$1; shift
    timeout 300 python bench.py --workload P4 --samples $S $B "$@" > gpurun_out/r2_c29_P4_${S}_$tag.log 2>&1
    echo "P4 $S $tag $(grep -o '"launch_ms": [0-9.]*' gpurun_out/r2_c29_P4_${S}_$tag.log | head -1) $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r2_c29_P4_${S}_$tag.log) $(grep -o '"verified_decryptions": [a-z]*' gpurun_out/r2_c29_P4_${S}_$tag.log)"
  done
done
