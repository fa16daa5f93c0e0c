#!/bin/bash
# GPU session 31 (1 GPU): final tree of round 1 -- full GPU suite, smoke, default bench (+cpu baseline), reference arm, launch list
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r31_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r31_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r31_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r31_smoke.log
timeout 600 python bench.py > gpurun_out/r31_bench_c2_gautschi.log 2>&1
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r31_bench_reference.log 2>&1
timeout 300 python bench.py --workload c1 --no-cpu > gpurun_out/r31_bench_c1.log 2>&1
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r31_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/r31_launches_c2_gautschi.csv python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r31_ncu.log 2>&1
grep -H -o '"ms_per_step": [0-9.]*' gpurun_out/r31_bench_*.log
