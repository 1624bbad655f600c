#!/bin/bash
# 1-GPU round-3 evidence run (gpurun -- 'bash tools/gpu_round3_n1.sh').  Every step is bounded; outputs under gpurun_out/r03_*.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r03_gpu_tests_n1_final.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/r03_gpu_tests_n1_final.log | cut -c1-300
timeout 900 python bench.py > gpurun_out/r03_bench_n1.json 2> gpurun_out/r03_bench_n1.err
echo "bench rc=$?"; tail -c 3000 gpurun_out/r03_bench_n1.json; tail -3 gpurun_out/r03_bench_n1.err | cut -c1-300
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r03_bench_ref_n1.json 2> gpurun_out/r03_bench_ref_n1.err
echo "ref rc=$?"; tail -c 800 gpurun_out/r03_bench_ref_n1.json
FLAGS="--steps 2 --warmup 3 --no-cpu --no-tp --no-flash --no-long-ctx --no-tp-e2e --no-fused-allreduce"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r03_launches_bench.csv \
    python bench.py $FLAGS > gpurun_out/r03_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:paged_decode -s 40 -c 2 -o gpurun_out/r03_decode_full -f \
    python bench.py $FLAGS --no-prefill --no-e2e > gpurun_out/r03_ncu_decode.log 2>&1
echo "ncu decode rc=$?"
ncu -i gpurun_out/r03_decode_full.ncu-rep --page raw --csv > gpurun_out/r03_decode_ncu_full_metrics.csv 2>/dev/null
echo "export rc=$?"; wc -c gpurun_out/r03_decode_ncu_full_metrics.csv
ls -la gpurun_out | tail -12
