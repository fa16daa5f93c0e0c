#!/bin/bash
# GPU session 28 (1 GPU): merged phase-A publish vs previous build; full GPU suite on the new build
set -x
mkdir -p gpurun_out
V=$PWD/nonlinear-solvers_b200/variants
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r28_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r28_pytest_gpu.log
for rep in 1 2; do
  timeout 300 python bench.py --steps 500 --warmup 20 --no-cpu > gpurun_out/r28_bench_c2_gautschi_merged_$rep.log 2>&1
  NLS_B200_LIB=$V/libnls_base.so timeout 300 python bench.py --steps 500 --warmup 20 --no-cpu > gpurun_out/r28_bench_c2_gautschi_base_$rep.log 2>&1
done
timeout 300 python bench.py --workload c1 --steps 500 --warmup 20 --no-cpu > gpurun_out/r28_bench_c1_merged.log 2>&1
NLS_B200_LIB=$V/libnls_base.so timeout 300 python bench.py --workload c1 --steps 500 --warmup 20 --no-cpu > gpurun_out/r28_bench_c1_base.log 2>&1
grep -H -o '"ms_per_step": [0-9.]*' gpurun_out/r28_bench_*.log
