set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_wop.py tests/test_gpu_sdfg.py -m gpu -x -q > gpurun_out/r2_pytest_c19.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_c19.log
tail -6 gpurun_out/r2_pytest_c19.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_1gpu_b.log 2>&1; echo "bench rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/r2_bench_1gpu_b.log'):
    if l.startswith('{'):
        j=json.loads(l); e=j['e2e']
        print('value',round(j['value']),'e2e',round(e['value']),'two_call',round(e['two_call']['value']),'sdfg',round(e['sdfg']['value']))
        print(json.dumps(j['configs'].get('configs[4] CRT/WoP-PBS'))[:600])
PY
