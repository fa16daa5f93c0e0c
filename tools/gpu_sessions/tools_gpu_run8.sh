#!/bin/bash
# GPU session 8 (1 GPU): TMA passes in the benches, persistent SV kernel, parity suite
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r8_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r8_pytest_gpu.log
timeout 300 python bench.py --workload c3 --steps 50 --warmup 5 --no-cpu > gpurun_out/r8_bench_c3.log 2>&1
timeout 300 python bench.py --workload c2_sv --steps 2000 --warmup 50 --no-cpu > gpurun_out/r8_bench_c2_sv.log 2>&1
timeout 600 python bench.py --gpus 1 --workload c5 --steps 10 --warmup 3 --no-cpu > gpurun_out/r8_bench_c5_n1.log 2>&1
timeout 300 python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu > gpurun_out/r8_plain_c3.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 30 -c 40 --csv --log-file gpurun_out/r8_launches_c3.csv python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu > gpurun_out/r8_ncu_c3.log 2>&1
ls -la gpurun_out | tail -8
