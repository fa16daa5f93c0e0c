#!/bin/bash
# GPU session 11 (1 GPU): phase-timing trace of the persistent kernel; default bench back on the kept loops; c3 cap 4 vs 6
set -x
mkdir -p gpurun_out
NLS_B200_LIB=$PWD/nonlinear-solvers_b200/variants/libnls_trace.so timeout 300 python tools/pk_trace.py > gpurun_out/r11_pk_trace.txt 2>&1
timeout 300 python bench.py --steps 500 --warmup 20 --no-cpu > gpurun_out/r11_bench_c2_gautschi.log 2>&1
for cap in 4 6; do
  NLS_FIRST_CAP=$cap timeout 300 python bench.py --workload c3 --steps 40 --warmup 5 --no-cpu > gpurun_out/r11_bench_c3_cap$cap.log 2>&1
done
timeout 300 python bench.py --workload c4 --steps 40 --warmup 5 --no-cpu > gpurun_out/r11_bench_c4.log 2>&1
grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/r11_bench_*.log
wc -l gpurun_out/r11_pk_trace.txt
