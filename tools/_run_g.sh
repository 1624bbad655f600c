timeout 300 python -m pytest tests/test_gpu_prefill_tcgen05.py tests/test_gpu_parity.py -m gpu -x -q -k "prefill or prefix" --timeout 120 > gpurun_out/r02_prefill_pair_tests.log 2>&1; echo "tests rc=$?"
tail -4 gpurun_out/r02_prefill_pair_tests.log | cut -c1-300
timeout 200 python tools/sweep_prefill.py --impls tcgen05 > gpurun_out/r02_prefill_sweep_pair_heads.jsonl 2>&1; echo "heads rc=$?"
NVLLM_PREFILL_PAIR=rows timeout 200 python tools/sweep_prefill.py --impls tcgen05 > gpurun_out/r02_prefill_sweep_pair_rows.jsonl 2>&1; echo "rows rc=$?"
echo HEADS; cut -c1-160 gpurun_out/r02_prefill_sweep_pair_heads.jsonl; echo ROWS; cut -c1-160 gpurun_out/r02_prefill_sweep_pair_rows.jsonl
