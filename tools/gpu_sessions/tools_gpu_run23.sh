#!/bin/bash
# GPU session 23 (1 GPU): full validation of the current tree + bench lines for every workload + launch list
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r23_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r23_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r23_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r23_smoke.log
timeout 600 python bench.py > gpurun_out/r23_bench_c2_gautschi.log 2>&1
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r23_bench_reference.log 2>&1
for w in c1 c2_sv; do timeout 300 python bench.py --workload $w --no-cpu > gpurun_out/r23_bench_$w.log 2>&1; done
for w in c3 c4; do timeout 300 python bench.py --workload $w --steps 50 --warmup 5 --no-cpu > gpurun_out/r23_bench_$w.log 2>&1; done
timeout 600 python bench.py --workload c5 --steps 10 --warmup 3 --no-cpu > gpurun_out/r23_bench_c5_n1.log 2>&1
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r23_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/r23_launches_c2_gautschi.csv python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r23_ncu.log 2>&1
grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/r23_bench_*.log
