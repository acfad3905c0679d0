set -x
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/r2_bench_4gpu.log 2>&1; echo "bench rc=$?"
tail -c 400 gpurun_out/r2_bench_4gpu.log
python - <<'PY'
import json
for l in open('gpurun_out/r2_bench_4gpu.log'):
    if l.startswith('{'):
        j=json.loads(l); e=j['e2e']
        print('N=4 value',round(j['value']),'ms',round(j['ms_per_step'],1),'e2e',round(e['value']),'two',round(e['two_call']['value']),'sdfg',round(e['sdfg']['value']),'fused',round(e['fused_extension']['value']))
        print('inlib', j['in_library_multi_gpu'])
        print('cfg3', {k:v for k,v in j['configs']['configs[3]'].items() if k in ('value','ms_per_step','n_gpus','tlus_this_gpu','verified_decryptions')})
PY
