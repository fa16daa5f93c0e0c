#!/bin/bash
# GPU session 40 (1 GPU): bench lines of every workload on the final library of the round + reference arm
set -x
mkdir -p gpurun_out
for w in c1 c2_sv c3 c4 c5; do
  timeout 400 python bench.py --workload $w --no-cpu > gpurun_out/r40_bench_$w.log 2>&1; echo "rc=$?" >> gpurun_out/r40_bench_$w.log
done
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r40_bench_reference.log 2>&1
grep -H -o '"ms_per_step": [0-9.]*\|"frac": [0-9.]*' gpurun_out/r40_bench_*.log
