#!/bin/bash
# 2-GPU validation of the fused RowParallelLinear (run through gpurun --gpus 2).  Every step is bounded.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/rp_gpus.txt 2>&1
nvidia-smi topo -m >> gpurun_out/rp_gpus.txt 2>&1
timeout 170 python -m pytest tests/test_gpu_rowparallel.py -q -m gpu --timeout 120 --maxfail 4 > gpurun_out/rp_tests.log 2>&1
echo "rp_tests rc=$?"; tail -30 gpurun_out/rp_tests.log | cut -c1-300
timeout 110 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29555 \
    tools/bench_rowparallel.py --iters 10 --out gpurun_out/rowparallel_tp2.jsonl > gpurun_out/rp_bench.log 2>&1
echo "rp_bench rc=$?"; tail -32 gpurun_out/rp_bench.log | cut -c1-420
NVLLM_FUSED_ALLREDUCE=1 timeout 150 python tools/bench_my_b200.py --attn-backend cuda --tensor-parallel-size 2 \
    --batch-sizes 16 64 --input-lens 512 --out gpurun_out/bench_my_tp2_cuda_fused_allreduce.json > gpurun_out/rp_bench_my_fused.log 2>&1
echo "bench_my fused rc=$?"; tail -6 gpurun_out/rp_bench_my_fused.log | cut -c1-300
NVLLM_FUSED_ALLREDUCE=0 timeout 150 python tools/bench_my_b200.py --attn-backend cuda --tensor-parallel-size 2 \
    --batch-sizes 16 64 --input-lens 512 --out gpurun_out/bench_my_tp2_cuda_nccl_allreduce.json > gpurun_out/rp_bench_my_nccl.log 2>&1
echo "bench_my nccl rc=$?"; tail -6 gpurun_out/rp_bench_my_nccl.log | cut -c1-300
