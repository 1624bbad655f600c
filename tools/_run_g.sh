timeout 300 python -m pytest tests/test_gpu_rowparallel_local.py -m gpu -x -q --timeout 120 > gpurun_out/r02_rp_tc_tests.log 2>&1; echo "tests rc=$?"
tail -15 gpurun_out/r02_rp_tc_tests.log | cut -c1-300
for tile in 64 128; do
  timeout 200 python tools/bench_rowparallel_local.py --tile $tile --tokens 64 128 256 512 --only 8b --iters 5 > gpurun_out/r02_rowparallel_local_tile${tile}.jsonl 2>&1; echo "bench tile $tile rc=$?"
  cut -c1-420 gpurun_out/r02_rowparallel_local_tile${tile}.jsonl | tail -14
done
