set -x
# v4 without TwT (P8): twiddles fetched one stage ahead (pf) / warp-uniform pass from the constant bank (c);
# 32-points-per-thread variant (--v4-alt 2).  Parity subset on the new default, A/B benches, fresh ncu.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "v4 or large_poly or cluster or external_product or fourier" > gpurun_out/r2_c28_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2_c28_pytest.log
tail -3 gpurun_out/r2_c28_pytest.log
B="--steps 2 --warmup 2 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
cp concrete_b200/libconcrete_b200.so /tmp/lib_main.so
for v in pf1c1 pf0c0 pf1c0 pf0c1; do
  cp build/ab/libconcrete_b200_$v.so concrete_b200/libconcrete_b200.so
  timeout 400 python bench.py --workload P8 --samples 1024 $B > gpurun_out/r2_c28_${v}_P8.log 2>&1
  echo "$v P8 1024 $(grep -o '"launch_ms": [0-9.]*' gpurun_out/r2_c28_${v}_P8.log) $(grep -o '"verified_decryptions": [a-z]*' gpurun_out/r2_c28_${v}_P8.log)"
done
for v in pf1c1 pf0c0; do
  cp build/ab/libconcrete_b200_$v.so concrete_b200/libconcrete_b200.so
  timeout 400 python bench.py --workload P8 --samples 1024 --v4-alt 2 $B > gpurun_out/r2_c28_${v}_P8_alt2.log 2>&1
  echo "$v P8 1024 alt2 $(grep -o '"launch_ms": [0-9.]*' gpurun_out/r2_c28_${v}_P8_alt2.log) $(grep -o '"verified_decryptions": [a-z]*' gpurun_out/r2_c28_${v}_P8_alt2.log)"
done
cp /tmp/lib_main.so concrete_b200/libconcrete_b200.so
A="--steps 1 --warmup 1 --no-configs --no-inlib --no-cpu-baseline --no-e2e --no-ref-gpu"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pbs_kernel -s 3 -c 1 -o gpurun_out/r2_c28_P8 python bench.py --workload P8 --samples 74 $A > gpurun_out/r2_c28_ncu_P8.log 2>&1
bash tools/ncu_export.sh gpurun_out/r2_c28_P8
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pbs_kernel -s 3 -c 1 -o gpurun_out/r2_c28_P8_alt2 python bench.py --workload P8 --samples 74 --v4-alt 2 $A > gpurun_out/r2_c28_ncu_P8_alt2.log 2>&1
bash tools/ncu_export.sh gpurun_out/r2_c28_P8_alt2
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pbs_kernel -s 3 -c 1 -o gpurun_out/r2_c28_P7 python bench.py --workload P7 --samples 148 $A > gpurun_out/r2_c28_ncu_P7.log 2>&1
bash tools/ncu_export.sh gpurun_out/r2_c28_P7
rm -f gpurun_out/*.source.csv.gz
ls -la gpurun_out | head -40
