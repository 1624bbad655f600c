timeout 200 python tools/sweep_decode.py --shapes c4 --variants -1 --splits 0,1,2,4,8,16 --clusters 0 > gpurun_out/r02_decode_sweep_c4.jsonl 2>&1; echo "c4 rc=$?"
timeout 200 python tools/sweep_decode.py --shapes c5all --variants -1 --flash --clusters 0,1 > gpurun_out/r02_decode_sweep_c5.jsonl 2>&1; echo "c5 rc=$?"
timeout 200 python tools/sweep_prefill.py --impls tcgen05 --flash > gpurun_out/r02_prefill_sweep.jsonl 2>&1; echo "prefill rc=$?"
cut -c1-260 gpurun_out/r02_decode_sweep_c4.jsonl | tail -30
