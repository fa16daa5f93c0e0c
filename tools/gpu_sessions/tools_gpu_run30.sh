#!/bin/bash
# GPU session 30 (1 GPU): CTA size of the persistent kernels (registers per thread vs warps)
set -x
mkdir -p gpurun_out
V=$PWD/nonlinear-solvers_b200/variants
timeout 300 python bench.py --steps 500 --warmup 20 --no-cpu > gpurun_out/r30_bench_c2_gautschi_t512.log 2>&1
for v in t256 t384 t640; do
  NLS_B200_LIB=$V/libnls_$v.so timeout 300 python bench.py --steps 500 --warmup 20 --no-cpu > gpurun_out/r30_bench_c2_gautschi_$v.log 2>&1
  NLS_B200_LIB=$V/libnls_$v.so timeout 300 python bench.py --workload c1 --steps 500 --warmup 20 --no-cpu > gpurun_out/r30_bench_c1_$v.log 2>&1
  NLS_B200_LIB=$V/libnls_$v.so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q > gpurun_out/r30_pytest_$v.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r30_pytest_$v.log
done
grep -H -o '"ms_per_step": [0-9.]*' gpurun_out/r30_bench_*.log
