set -x
# final evidence of the round on one B200: full GPU suite, smoke(), the bench line, the reference arm, the config sweep with
# the reference GPU library beside, the launch list and fresh ncu captures of the kernels that changed last (P4, P6)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_final4_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_final4_pytest.log
tail -3 gpurun_out/r2_final4_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_final4_smoke.log 2>&1; tail -1 gpurun_out/r2_final4_smoke.log
timeout 1200 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_final4_bench.log 2>&1; tail -c 600 gpurun_out/r2_final4_bench.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_final4_reference_arm.log 2>&1; tail -c 400 gpurun_out/r2_final4_reference_arm.log
B="--steps 3 --warmup 3 --no-configs --no-inlib --no-cpu-baseline --no-e2e"
: > gpurun_out/r2_final4_sweep.jsonl
for W in "P4 16384" "P4 64" "P6 4096" "P7 4096" "P8 1024"; do set -- $W
  timeout 900 python bench.py --workload $1 --samples $2 $B 2>/dev/null | grep '^{' >> gpurun_out/r2_final4_sweep.jsonl
done
timeout 600 python bench.py --workload P8 --samples 4096 --steps 2 --warmup 2 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu 2>/dev/null | grep '^{' >> gpurun_out/r2_final4_sweep.jsonl
wc -l gpurun_out/r2_final4_sweep.jsonl
A="--no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_final4_launches.csv python bench.py --steps 2 --warmup 3 $A > gpurun_out/r2_final4_ncu_launches.log 2>&1
rm -f gpurun_out/*.source.csv.gz
