#!/bin/bash
# GPU session 4 (1 GPU): full parity suite after the Gram-corrected 2-pass Lanczos, benches, c3 launch list
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r4_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r4_pytest_gpu.log
for w in c3 c4 c2_sv; do
  timeout 300 python bench.py --workload $w --steps 100 --warmup 5 --no-cpu > gpurun_out/r4_bench_$w.log 2>&1
done
timeout 600 python bench.py --gpus 1 --workload c5 --steps 10 --warmup 3 --no-cpu > gpurun_out/r4_bench_c5_n1.log 2>&1
timeout 300 python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu > gpurun_out/r4_plain_c3.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 30 -c 60 --csv --log-file gpurun_out/r4_launches_c3.csv python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu > gpurun_out/r4_ncu_c3.log 2>&1
ls -la gpurun_out | tail -12
