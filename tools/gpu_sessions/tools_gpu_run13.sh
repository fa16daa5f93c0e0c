#!/bin/bash
# GPU session 13 (1 GPU): register-carried QL chain (default build) + A/B of CTA size / column-group variants
set -x
mkdir -p gpurun_out
V=$PWD/nonlinear-solvers_b200/variants
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r13_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r13_pytest_gpu.log
for w in c2_gautschi c1; do
  timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu > gpurun_out/r13_bench_${w}_base.log 2>&1
  for v in t1024_nv4 t1024_nv8 t512_nv4; do
    NLS_B200_LIB=$V/libnls_$v.so timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu > gpurun_out/r13_bench_${w}_$v.log 2>&1
  done
done
NLS_B200_LIB=$V/libnls_t1024_nv4.so timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r13_pytest_t1024.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r13_pytest_t1024.log
NLS_B200_LIB=$V/libnls_trace.so timeout 300 python tools/pk_trace.py > gpurun_out/r13_pk_trace.txt 2>&1
grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/r13_bench_*.log
