#!/bin/bash
# GPU session 1: parity suite, smoke, all bench workloads, ncu launch list + full capture of the persistent kernel
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r1_gpus.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r1_smoke.log 2>&1
timeout 300 python bench.py > gpurun_out/r1_bench_c2_gautschi.log 2>&1
for w in c1 c2_sv c3 c4; do
  timeout 300 python bench.py --workload $w --steps 100 --warmup 5 --no-cpu > gpurun_out/r1_bench_$w.log 2>&1
done
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1_bench_reference.log 2>&1
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/r1_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r1_launches_c2_gautschi.csv python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/r1_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_pk_gautschi -s 1 -c 1 -o gpurun_out/r1_pk_gautschi python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/r1_ncu2.log 2>&1
ls -la gpurun_out
