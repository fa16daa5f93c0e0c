#!/bin/bash
# GPU session 20: address-space hints local to the site-loop functions (A, B, Gautschi mid/tail, all)
set -x
mkdir -p gpurun_out
V=$PWD/nonlinear-solvers_b200/variants
for v in ha hb hg habg; do
  NLS_B200_LIB=$V/libnls_$v.so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q > gpurun_out/r20_pytest_$v.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r20_pytest_$v.log
  for w in c2_gautschi c1; do
    NLS_B200_LIB=$V/libnls_$v.so timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu > gpurun_out/r20_bench_${w}_$v.log 2>&1
  done
done
grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/r20_bench_*.log
