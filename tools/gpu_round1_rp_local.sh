#!/bin/bash
# 1-GPU tuning run for the GEMM part of the fused RowParallelLinear kernel (exchange off).  Every step is bounded.
mkdir -p gpurun_out
timeout 150 python tools/bench_rowparallel_local.py --out gpurun_out/rowparallel_local_bk128.jsonl > gpurun_out/rp_local_bk128.log 2>&1
echo "bk auto rc=$?"; tail -22 gpurun_out/rp_local_bk128.log | cut -c1-420
NVLLM_RP_BK=64 timeout 150 python tools/bench_rowparallel_local.py --out gpurun_out/rowparallel_local_bk64.jsonl > gpurun_out/rp_local_bk64.log 2>&1
echo "bk64 rc=$?"; tail -22 gpurun_out/rp_local_bk64.log | cut -c1-420
timeout 150 ncu --set full --clock-control none --import-source on -k regex:rowparallel -s 3 -c 1 -f -o gpurun_out/rp_local_gemm \
    python tools/bench_rowparallel_local.py --ncu > gpurun_out/rp_local_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/rp_local_ncu.log
