set -x
# N = 16384: 32 points per thread is the default now.  Full GPU suite, P8 at 1024 / 4096 / 74, ncu of the new default.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c38_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_c38_pytest.log
tail -3 gpurun_out/r2_c38_pytest.log
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e"
: > gpurun_out/r2_c38_sweep.jsonl
for W in "P8 1024" "P8 74"; do set -- $W
  timeout 900 python bench.py --workload $1 --samples $2 $B 2>/dev/null | grep '^{' >> gpurun_out/r2_c38_sweep.jsonl
done
timeout 600 python bench.py --workload P8 --samples 4096 --steps 2 --warmup 2 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu 2>/dev/null | grep '^{' >> gpurun_out/r2_c38_sweep.jsonl
timeout 600 python bench.py --workload P4 --samples 16384 $B 2>/dev/null | grep '^{' >> gpurun_out/r2_c38_sweep.jsonl
python - <<'PY'
import json
for l in open('gpurun_out/r2_c38_sweep.jsonl'):
    j=json.loads(l); print(j['config']['workload'][36:80], round(j['roofline']['launch_ms'],2), round(j['roofline']['frac'],4), j['verified_decryptions'], j['vs_reference_gpu'])
PY
A="--steps 1 --warmup 1 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pbs_kernel -s 3 -c 1 -o gpurun_out/r2_c38_P8 python bench.py --workload P8 --samples 74 $A > gpurun_out/r2_c38_ncu_P8.log 2>&1
bash tools/ncu_export.sh gpurun_out/r2_c38_P8
rm -f gpurun_out/*.source.csv.gz
