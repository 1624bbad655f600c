timeout 300 python -m pytest tests/test_gpu_decode_state.py tests/test_gpu_engine_e2e.py tests/test_gpu_sampler_sharded.py tests/test_gpu_sampler.py tests/test_gpu_multi_device.py -m gpu -x -q --timeout 200 > gpurun_out/r02_decode_state_tests.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/r02_decode_state_tests.log | cut -c1-300
