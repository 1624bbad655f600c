#!/bin/bash
# N-GPU validation + measurement (gpurun --gpus N -- 'bash tools/gpu_round2_tp.sh N').  Every step is bounded.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/r02_tp${N}_gpus.txt 2>&1
nvidia-smi topo -m >> gpurun_out/r02_tp${N}_gpus.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_rowparallel.py tests/test_gpu_rowparallel_worlds.py tests/test_gpu_multi_device.py -q -m gpu \
    --timeout 150 --maxfail 6 > gpurun_out/r02_tp${N}_tests.log 2>&1
echo "tests rc=$?"; tail -12 gpurun_out/r02_tp${N}_tests.log | cut -c1-300
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29555 \
    tools/bench_rowparallel.py --iters 10 --out gpurun_out/r02_rowparallel_tp${N}.jsonl > gpurun_out/r02_tp${N}_rp_bench.log 2>&1
echo "rp_bench rc=$?"; tail -24 gpurun_out/r02_tp${N}_rp_bench.log | cut -c1-400
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29566 \
    bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02_bench_n${N}.json 2> gpurun_out/r02_bench_n${N}.err
echo "bench rc=$?"; tail -c 6000 gpurun_out/r02_bench_n${N}.json; tail -5 gpurun_out/r02_bench_n${N}.err | cut -c1-300
for fused in 1 0; do
  NVLLM_FUSED_ALLREDUCE=$fused timeout 240 python tools/bench_my_b200.py --attn-backend cuda --tensor-parallel-size $N \
      --batch-sizes 16 64 --input-lens 512 --port $((2400 + fused * 10)) --out gpurun_out/r02_bench_my_tp${N}_fused${fused}.json \
      > gpurun_out/r02_tp${N}_bench_my_fused${fused}.log 2>&1
  echo "bench_my fused=$fused rc=$?"; tail -4 gpurun_out/r02_tp${N}_bench_my_fused${fused}.log | cut -c1-300
done
