# N-GPU bench.py collective legs only (gpurun --gpus N -- bash tools/gpu_round2_bench_only.sh N).
N=${1:-8}
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29588 \
    bench.py --gpus $N --steps 5 --warmup 3 --no-tp-e2e --no-long-ctx --no-prefill --no-flash --no-e2e > gpurun_out/r02_bench_n${N}_tc.json 2> gpurun_out/r02_bench_n${N}_tc.err
echo "bench rc=$?"; tail -3 gpurun_out/r02_bench_n${N}_tc.err | cut -c1-300
