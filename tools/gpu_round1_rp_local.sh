#!/bin/bash
# 1-GPU tuning run for the GEMM part of the fused RowParallelLinear kernel (exchange off).  Every step is bounded.
mkdir -p gpurun_out
timeout 100 python tools/bench_rowparallel_local.py --iters 4 --out gpurun_out/rowparallel_local.jsonl > gpurun_out/rp_local.log 2>&1
echo "local rc=$?"; python - <<'PY'
import json
for l in open('gpurun_out/rowparallel_local.jsonl'):
    r = json.loads(l)
    print(r['shape'], r['n'], r['plan'], 'diff', r['max_abs_diff_vs_cublas'], 'fused', r['us_fused_hbm'], 'cublas', r['us_cublas_hbm'], 'GB/s', r['w_gbs_fused_hbm'])
PY
timeout 60 ncu --set full --clock-control none --import-source on -k regex:rowparallel -s 3 -c 1 -f -o gpurun_out/rp_local_gemm_v5 \
    python tools/bench_rowparallel_local.py --ncu > gpurun_out/rp_local_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/rp_local_ncu.log
