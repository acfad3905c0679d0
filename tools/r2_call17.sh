set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sdfg.py tests/test_gpu_wop.py -m gpu -x -q > gpurun_out/r2_pytest_c17.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_c17.log
tail -4 gpurun_out/r2_pytest_c17.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-configs > gpurun_out/r2_c17_e2e.log 2>&1
python - <<'PY'
import json
for l in open('gpurun_out/r2_c17_e2e.log'):
    if l.startswith('{'):
        e=json.loads(l)['e2e']; print('e2e',round(e['value']),round(e['ms_per_step'],1),'two_call',round(e['two_call']['value']),round(e['two_call']['ms_per_step'],1),'sdfg',round(e['sdfg']['value']),round(e['sdfg']['ms_per_step'],1))
PY
