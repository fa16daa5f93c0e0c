#!/bin/bash
# GPU session 29 (1 GPU): previous-column shared tile in the persistent Lanczos kernels
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r29_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r29_pytest_gpu.log
for rep in 1 2; do
  timeout 300 python bench.py --steps 500 --warmup 20 --no-cpu > gpurun_out/r29_bench_c2_gautschi_xprev_$rep.log 2>&1
  NLS_BENCH_FLAGS=128 timeout 300 python bench.py --steps 500 --warmup 20 --no-cpu > gpurun_out/r29_bench_c2_gautschi_noxprev_$rep.log 2>&1
done
timeout 300 python bench.py --workload c1 --steps 500 --warmup 20 --no-cpu > gpurun_out/r29_bench_c1_xprev.log 2>&1
NLS_BENCH_FLAGS=128 timeout 300 python bench.py --workload c1 --steps 500 --warmup 20 --no-cpu > gpurun_out/r29_bench_c1_noxprev.log 2>&1
grep -H -o '"ms_per_step": [0-9.]*' gpurun_out/r29_bench_*.log
