#!/bin/bash
# GPU session 15: bisect the address-space hints (which assumption is false in k_pk_gautschi)
set -x
mkdir -p gpurun_out
V=$PWD/nonlinear-solvers_b200/variants
for v in noglobal noshared; do
  NLS_B200_LIB=$V/libnls_$v.so timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "test_real_2d_steps" > gpurun_out/r15_pytest_$v.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r15_pytest_$v.log
  NLS_B200_LIB=$V/libnls_$v.so timeout 300 python bench.py --steps 500 --warmup 20 --no-cpu > gpurun_out/r15_bench_c2_gautschi_$v.log 2>&1
done
tail -3 gpurun_out/r15_pytest_*.log
