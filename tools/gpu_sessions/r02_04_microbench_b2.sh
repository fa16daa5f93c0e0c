#!/bin/bash
# GPU session r02-04 (1 GPU): micro-benchmarks of the grid-wide all-reduce schemes and of the tridiagonal
# f(T)e1 solvers (tools/microbench), and the B2 class-layer tests.
set -x
mkdir -p gpurun_out
timeout 300 ./tools/microbench/allreduce_bench > gpurun_out/r02d_allreduce_bench.log 2>&1; echo "rc=$?" >> gpurun_out/r02d_allreduce_bench.log
cat gpurun_out/r02d_allreduce_bench.log
timeout 300 ./tools/microbench/tridiag_bench > gpurun_out/r02d_tridiag_bench.log 2>&1; echo "rc=$?" >> gpurun_out/r02d_tridiag_bench.log
cat gpurun_out/r02d_tridiag_bench.log
( time timeout 900 python -m pytest tests/test_b2_classes.py -m gpu -q -rA ) > gpurun_out/r02d_pytest_b2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02d_pytest_b2.log
grep -E "passed|failed|vs CLI|vs oracle|rc=|Error|error" gpurun_out/r02d_pytest_b2.log | head -40
