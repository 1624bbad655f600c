#!/bin/bash
# N-GPU validation + measurement (gpurun --gpus N -- 'bash tools/gpu_round2_tp.sh N [lean]').  Every step is bounded.
N=${1:-2}
LEAN=${2:-}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/r02_tp${N}_gpus.txt 2>&1
nvidia-smi topo -m >> gpurun_out/r02_tp${N}_gpus.txt 2>&1
timeout 700 python -m pytest tests/test_gpu_rowparallel.py tests/test_gpu_rowparallel_worlds.py tests/test_gpu_multi_device.py \
    tests/test_gpu_sampler_sharded.py tests/test_gpu_vocab_embedding.py tests/test_gpu_engine_tp.py -q -m gpu --timeout 300 --maxfail 6 > gpurun_out/r02_tp${N}_tests.log 2>&1
echo "tests rc=$?"; tail -12 gpurun_out/r02_tp${N}_tests.log | cut -c1-300
if [ -z "$LEAN" ]; then
  timeout 600 python tools/check_tp_tokens.py --tensor-parallel-size $N --layers 28 --steps 64 --port 2480 \
      --out gpurun_out/r02_tp${N}_tokens.json > gpurun_out/r02_tp${N}_tokens.log 2>&1
  echo "tokens rc=$?"; tail -2 gpurun_out/r02_tp${N}_tokens.log | cut -c1-1200
  EXTRA=""
else
  EXTRA="--no-tp-e2e --no-long-ctx --no-prefill --no-flash"
fi
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29566 \
    bench.py --gpus $N --steps 5 --warmup 3 $EXTRA > gpurun_out/r02_bench_n${N}.json 2> gpurun_out/r02_bench_n${N}.err
echo "bench rc=$?"; tail -c 9000 gpurun_out/r02_bench_n${N}.json; tail -5 gpurun_out/r02_bench_n${N}.err | cut -c1-300
