#!/bin/bash
# GPU session 14 (1 GPU): persistent kernels with address-space hints (LDS/STS instead of generic LD/ST)
set -x
mkdir -p gpurun_out
V=$PWD/nonlinear-solvers_b200/variants
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r14_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r14_pytest_gpu.log
for w in c2_gautschi c1 c2_sv; do
  timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu > gpurun_out/r14_bench_${w}.log 2>&1
done
NLS_B200_LIB=$V/libnls_trace.so timeout 300 python tools/pk_trace.py > gpurun_out/r14_pk_trace.txt 2>&1
grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/r14_bench_*.log
