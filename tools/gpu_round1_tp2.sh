#!/bin/bash
# 2-GPU validation of the fused RowParallelLinear (run through gpurun --gpus 2).  Every step is bounded.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/rp_gpus.txt 2>&1
nvidia-smi topo -m >> gpurun_out/rp_gpus.txt 2>&1
timeout 420 python -m pytest tests/test_gpu_rowparallel.py -q -m gpu --timeout 150 > gpurun_out/rp_tests.log 2>&1
echo "rp_tests rc=$?"; tail -25 gpurun_out/rp_tests.log
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29555 \
    tools/bench_rowparallel.py --out gpurun_out/rowparallel_tp2.jsonl > gpurun_out/rp_bench.log 2>&1
echo "rp_bench rc=$?"; tail -12 gpurun_out/rp_bench.log | cut -c1-400
NVLLM_FUSED_ALLREDUCE=1 timeout 330 python tools/bench_my_b200.py --attn-backend cuda --tensor-parallel-size 2 \
    --batch-sizes 16 64 --input-lens 512 --out gpurun_out/bench_my_tp2_cuda_fused_allreduce.json > gpurun_out/rp_bench_my.log 2>&1
echo "bench_my rc=$?"; tail -8 gpurun_out/rp_bench_my.log | cut -c1-300
