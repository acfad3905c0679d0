set -x
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 10 --warmup 5 > gpurun_out/r2_bench_8gpu.log 2>&1; echo "bench rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/r2_bench_8gpu.log'):
    if l.startswith('{'):
        j=json.loads(l); e=j['e2e']
        print('N=8 value',round(j['value']),'ms',round(j['ms_per_step'],1),'e2e',round(e['value']),'two',round(e['two_call']['value']),'sdfg',round(e['sdfg']['value']),'fused',round(e['fused_extension']['value']))
        print('inlib', {k:v for k,v in j['in_library_multi_gpu'].items() if k!='api'})
        print('cfg3', {k:v for k,v in j['configs']['configs[3]'].items() if k in ('value','ms_per_step','n_gpus','tlus_this_gpu','verified_decryptions')})
PY
tail -c 300 gpurun_out/r2_bench_8gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29532 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 2>&1 | grep '^{' | cut -c1-300
