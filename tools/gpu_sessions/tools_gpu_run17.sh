#!/bin/bash
# GPU session 17: address-space hints with the pointer/integer round trip removed
set -x
mkdir -p gpurun_out
V=$PWD/nonlinear-solvers_b200/variants
NLS_B200_LIB=$V/libnls_hints.so timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q > gpurun_out/r17_pytest_hints.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r17_pytest_hints.log
for w in c2_gautschi c1; do
  NLS_B200_LIB=$V/libnls_hints.so timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu > gpurun_out/r17_bench_${w}_hints.log 2>&1
  timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu > gpurun_out/r17_bench_${w}_base.log 2>&1
done
NLS_B200_LIB=$V/libnls_trace.so timeout 300 python tools/pk_trace.py > gpurun_out/r17_pk_trace.txt 2>&1
grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/r17_bench_*.log
