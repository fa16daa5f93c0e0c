#!/bin/bash
# GPU session 12 (1 GPU): persistent kernel with batched gather/halo loads, merged latency window, rsqrt QL;
# A/B of loop-unroll variants; parity on the default build and on the unroll variant (pk tests)
set -x
mkdir -p gpurun_out
V=$PWD/nonlinear-solvers_b200/variants
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r12_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r12_pytest_gpu.log
for w in c2_gautschi c1; do
  timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu > gpurun_out/r12_bench_${w}_base.log 2>&1
  for v in unroll comb4; do
    NLS_B200_LIB=$V/libnls_$v.so timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu > gpurun_out/r12_bench_${w}_$v.log 2>&1
  done
done
NLS_B200_LIB=$V/libnls_unroll.so timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r12_pytest_unroll.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r12_pytest_unroll.log
NLS_B200_LIB=$V/libnls_trace.so timeout 300 python tools/pk_trace.py > gpurun_out/r12_pk_trace.txt 2>&1
grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/r12_bench_*.log
