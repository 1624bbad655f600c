#!/bin/bash
# 2-GPU validation + measurement of the fused RowParallelLinear (gpurun --gpus 2).  Every step is bounded.
mkdir -p gpurun_out
timeout 100 python tools/bench_rowparallel_local.py --iters 5 --out gpurun_out/rowparallel_local.jsonl > gpurun_out/rp_local.log 2>&1
echo "local rc=$?"; python - <<'PY'
import json
for l in open('gpurun_out/rowparallel_local.jsonl'):
    r = json.loads(l)
    print(r['shape'], r['n'], r['plan'], 'diff', r['max_abs_diff_vs_cublas'], 'fused', r['us_fused_hbm'], 'cublas', r['us_cublas_hbm'], 'GB/s', r['w_gbs_fused_hbm'])
PY
bash tools/gpu_round1_tp2.sh
