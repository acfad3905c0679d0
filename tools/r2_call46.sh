set -x
# 8-GPU box, final kernels: torchrun bench line (weak scaling on configs[1], configs[3] sharded, in-library one-process leg)
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2_final_bench_8gpu.log 2>&1; echo "bench rc=$?"
tail -c 1200 gpurun_out/r2_final_bench_8gpu.log
