set -x
# 2-GPU box, final kernels: the two multi-GPU tests that are skipped on 1-GPU boxes, torchrun bench line (with configs[3] sharded and the in-library leg)
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests -m gpu -x -q -k "two_gpus or sharded or multi_gpu" > gpurun_out/r2_final_pytest_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_final_pytest_2gpu.log
tail -4 gpurun_out/r2_final_pytest_2gpu.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_final_bench_2gpu.log 2>&1; echo "bench rc=$?"
tail -c 1500 gpurun_out/r2_final_bench_2gpu.log
