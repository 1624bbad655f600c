timeout 300 python -m pytest tests/test_gpu_rowparallel_local.py -m gpu -x -q --timeout 120 -k tcgen05 > gpurun_out/r02_rp_tc_tests.log 2>&1; echo "tests rc=$?"
tail -4 gpurun_out/r02_rp_tc_tests.log | cut -c1-300
timeout 200 python tools/bench_rowparallel_local.py --tile 128 --tokens 64 128 256 512 --only 8b --iters 5 > gpurun_out/r02_rowparallel_local_tile128_v2.jsonl 2>&1; echo "bench rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/r02_rowparallel_local_tile128_v2.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print(d['shape'], d['n'], d['plan'], 'fused', d['us_fused_hbm'], 'cublas', d['us_cublas_hbm'])
PY
