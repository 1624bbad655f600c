# 2-GPU check of the tcgen05 RowParallelLinear tile: rowparallel tests + bench.py collective legs (gpurun --gpus 2).
timeout 500 python -m pytest tests/test_gpu_rowparallel.py tests/test_gpu_rowparallel_local.py -q -m gpu --timeout 200 --maxfail 5 > gpurun_out/r02_tp2_tc_tests.log 2>&1; echo "tests rc=$?"
tail -8 gpurun_out/r02_tp2_tc_tests.log | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29577 \
    bench.py --gpus 2 --steps 5 --warmup 3 --no-tp-e2e --no-long-ctx --no-prefill --no-flash > gpurun_out/r02_bench_n2_tc.json 2> gpurun_out/r02_bench_n2_tc.err
echo "bench rc=$?"; tail -3 gpurun_out/r02_bench_n2_tc.err | cut -c1-300
