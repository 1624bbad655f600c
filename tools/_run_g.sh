set -x
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r02_gpu_tests_h.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/r02_gpu_tests_h.log
timeout 600 python tools/sweep_decode.py --shapes c5all --variants=-1 --splits 0,8,16,32,64 --clusters 1,0,8 > gpurun_out/r02_decode_sweep_cluster_v1.jsonl 2>&1; echo "sweep rc=$?"
timeout 300 python tools/sweep_decode.py --shapes c1,c2 --variants=-1 --splits 0 --clusters 1,0 >> gpurun_out/r02_decode_sweep_cluster_v1.jsonl 2>&1
timeout 120 python tools/trace_decode.py --bs 1 --hkv 8 --hq 32 --splits 16 > gpurun_out/r02_decode_trace_v4.jsonl 2>&1
timeout 120 python tools/trace_decode.py --bs 1 --hkv 1 --hq 4 --splits 64 >> gpurun_out/r02_decode_trace_v4.jsonl 2>&1
timeout 120 python tools/trace_decode.py --bs 1 --hkv 8 --hq 32 --splits 16 --cluster 1 >> gpurun_out/r02_decode_trace_v4.jsonl 2>&1
