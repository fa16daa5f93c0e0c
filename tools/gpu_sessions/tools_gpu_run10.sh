#!/bin/bash
# GPU session 10 (1 GPU): A/B of the persistent kernel's two-stage load blocks (3 builds of the same tree)
# and of the complex pass-A first-chunk cap; parity suite on the default build first
set -x
mkdir -p gpurun_out
V=nonlinear-solvers_b200/variants
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r10_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r10_pytest_gpu.log
for v in v0_old v1_inline v2_noinline; do
  for w in c2_gautschi c1; do
    NLS_B200_LIB=$PWD/$V/libnls_$v.so timeout 300 python bench.py --workload $w --steps 500 --warmup 20 --no-cpu > gpurun_out/r10_bench_${w}_$v.log 2>&1
  done
done
for cap in 6 8 10 12; do
  NLS_FIRST_CAP=$cap timeout 300 python bench.py --workload c4 --steps 40 --warmup 5 --no-cpu > gpurun_out/r10_bench_c4_cap$cap.log 2>&1
  NLS_FIRST_CAP=$cap timeout 300 python bench.py --workload c3 --steps 40 --warmup 5 --no-cpu > gpurun_out/r10_bench_c3_cap$cap.log 2>&1
  NLS_FIRST_CAP=$cap timeout 600 python bench.py --workload c5 --steps 8 --warmup 3 --no-cpu > gpurun_out/r10_bench_c5_cap$cap.log 2>&1
done
grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/r10_bench_*.log | head -40
ls gpurun_out | grep r10 | wc -l
