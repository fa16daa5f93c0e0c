#!/bin/bash
# GPU session 25: runtime-NC site loops (instruction-cache residency) vs templated
set -x
mkdir -p gpurun_out
V=$PWD/nonlinear-solvers_b200/variants
NLS_B200_LIB=$V/libnls_rt.so timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q > gpurun_out/r25_pytest_rt.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r25_pytest_rt.log
for w in c2_gautschi c1; do
  NLS_B200_LIB=$V/libnls_rt.so timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu > gpurun_out/r25_bench_${w}_rt.log 2>&1
  timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu > gpurun_out/r25_bench_${w}_base.log 2>&1
done
grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/r25_bench_*.log
