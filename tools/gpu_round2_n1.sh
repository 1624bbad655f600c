#!/bin/bash
# 1-GPU round-2 evidence run (gpurun -- 'bash tools/gpu_round2_n1.sh').  Every step is bounded; outputs under gpurun_out/r02_*.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02_gpu_tests_final.log 2>&1
echo "tests rc=$?"; tail -6 gpurun_out/r02_gpu_tests_final.log | cut -c1-300
timeout 900 python bench.py > gpurun_out/r02_bench_n1_final.json 2> gpurun_out/r02_bench_n1_final.err
echo "bench rc=$?"; tail -c 7000 gpurun_out/r02_bench_n1_final.json; tail -3 gpurun_out/r02_bench_n1_final.err | cut -c1-300
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_ref_n1.json 2> gpurun_out/r02_bench_ref_n1.err
echo "ref rc=$?"; tail -c 1500 gpurun_out/r02_bench_ref_n1.json
FLAGS="--steps 2 --warmup 3 --no-cpu --no-tp --no-flash --no-long-ctx --no-tp-e2e --no-fused-allreduce"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_bench_final.csv \
    python bench.py $FLAGS > gpurun_out/r02_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:paged_decode -s 40 -c 2 -o gpurun_out/r02_decode_full_final -f \
    python bench.py $FLAGS --no-prefill --no-e2e > gpurun_out/r02_ncu_decode.log 2>&1
echo "ncu decode rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:prefill_tcgen05 -s 6 -c 2 -o gpurun_out/r02_prefill_full_final -f \
    python bench.py $FLAGS --no-e2e > gpurun_out/r02_ncu_prefill.log 2>&1
echo "ncu prefill rc=$?"
ls -la gpurun_out | tail -20
