#!/bin/bash
# GPU session 22: rewritten persistent Stoermer-Verlet kernel (in-place tiles, restrict loop) vs the per-step graph
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q > gpurun_out/r22_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r22_pytest_gpu.log
timeout 300 python bench.py --workload c2_sv --steps 2000 --warmup 50 --no-cpu > gpurun_out/r22_bench_c2_sv_graph.log 2>&1
NLS_BENCH_FLAGS=64 timeout 300 python bench.py --workload c2_sv --steps 2000 --warmup 50 --no-cpu > gpurun_out/r22_bench_c2_sv_pk.log 2>&1
timeout 300 python bench.py --steps 500 --warmup 20 --no-cpu > gpurun_out/r22_bench_c2_gautschi.log 2>&1
grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/r22_bench_*.log
