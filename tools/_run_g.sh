timeout 120 python -m pytest tests/test_gpu_prefill_tcgen05.py tests/test_gpu_parity.py -m gpu -x -q -k "prefill or prefix" > gpurun_out/r02_gpu_tests_o.log 2>&1; rc=$?; echo "tests rc=$rc"
tail -3 gpurun_out/r02_gpu_tests_o.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 200 python tools/sweep_prefill.py --impls tcgen05 --flash > gpurun_out/r02_prefill_sweep_c.jsonl 2>&1; echo "sweep rc=$?"
cut -c1-200 gpurun_out/r02_prefill_sweep_c.jsonl
