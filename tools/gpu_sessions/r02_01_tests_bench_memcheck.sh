#!/bin/bash
# GPU session r02-01 (1 GPU): full GPU suite incl. the new full-size parity tests, the new default bench line
# (C5 headline + "others"), the reference arm, and compute-sanitizer memcheck on the small kernel-family cases.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r02a_smi.txt
nproc >> gpurun_out/r02a_smi.txt; free -g >> gpurun_out/r02a_smi.txt
( time timeout 2400 python -m pytest tests -m gpu -q -rs --durations=25 ) > gpurun_out/r02a_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02a_pytest_gpu.log
tail -60 gpurun_out/r02a_pytest_gpu.log
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 ) > gpurun_out/r02a_bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/r02a_bench.log
tail -5 gpurun_out/r02a_bench.log
( time timeout 600 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 ) > gpurun_out/r02a_bench_reference.log 2>&1
tail -3 gpurun_out/r02a_bench_reference.log
timeout 300 python tools/sanitize_cases.py > gpurun_out/r02a_sanitize_plain.log 2>&1 && \
timeout 1500 compute-sanitizer --tool memcheck --print-limit 50 python tools/sanitize_cases.py > gpurun_out/r02a_memcheck.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/r02a_memcheck.log
tail -15 gpurun_out/r02a_memcheck.log
