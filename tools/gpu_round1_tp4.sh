#!/bin/bash
# 4-GPU measurement of the fused RowParallelLinear (gpurun --gpus 4).  Every step is bounded.
mkdir -p gpurun_out
timeout 110 python -m torch.distributed.run --nnodes=1 --nproc-per-node=4 --master-addr 127.0.0.1 --master-port 29556 \
    tools/bench_rowparallel.py --iters 10 --out gpurun_out/rowparallel_tp4.jsonl > gpurun_out/rp_bench_tp4.log 2>&1
echo "rp_bench tp4 rc=$?"; tail -3 gpurun_out/rp_bench_tp4.log | cut -c1-200
NVLLM_FUSED_ALLREDUCE=1 timeout 150 python tools/bench_my_b200.py --attn-backend cuda --tensor-parallel-size 4 \
    --batch-sizes 16 64 --input-lens 512 --out gpurun_out/bench_my_tp4_cuda_fused_allreduce.json > gpurun_out/rp_bench_my_fused_tp4.log 2>&1
echo "bench_my fused rc=$?"; tail -6 gpurun_out/rp_bench_my_fused_tp4.log | cut -c1-300
NVLLM_FUSED_ALLREDUCE=0 timeout 150 python tools/bench_my_b200.py --attn-backend cuda --tensor-parallel-size 4 \
    --batch-sizes 16 64 --input-lens 512 --out gpurun_out/bench_my_tp4_cuda_nccl_allreduce.json > gpurun_out/rp_bench_my_nccl_tp4.log 2>&1
echo "bench_my nccl rc=$?"; tail -6 gpurun_out/rp_bench_my_nccl_tp4.log | cut -c1-300
