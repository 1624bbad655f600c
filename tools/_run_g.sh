timeout 300 python -m pytest tests/test_gpu_vocab_embedding.py tests/test_gpu_engine_e2e.py -m gpu -x -q --timeout 200 > gpurun_out/r02_embedding_tests.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/r02_embedding_tests.log | cut -c1-300
