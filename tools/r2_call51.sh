set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_workloads.py -m gpu -x -q > gpurun_out/r2_c51_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_c51_pytest.log
tail -3 gpurun_out/r2_c51_pytest.log
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e"
: > gpurun_out/r2_c51_sweep.jsonl
for W in "P4 16384" "P4 4096" "P4 64"; do set -- $W
  timeout 900 python bench.py --workload $1 --samples $2 $B 2>/dev/null | grep '^{' >> gpurun_out/r2_c51_sweep.jsonl
done
python - <<'PY'
import json
for l in open('gpurun_out/r2_c51_sweep.jsonl'):
    j=json.loads(l); print(j['config']['workload'][36:80], round(j['roofline']['launch_ms'],2), round(j['roofline']['frac'],4), j['verified_decryptions'], j['vs_reference_gpu'], round(j['value']))
PY
