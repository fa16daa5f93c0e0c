#!/bin/bash
# GPU session 5 (2 GPUs): parity suite incl. sEWI, march pass, slabs; benches
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r5_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r5_pytest_gpu.log
timeout 300 python bench.py --workload c3 --steps 100 --warmup 5 --no-cpu > gpurun_out/r5_bench_c3.log 2>&1
timeout 600 python bench.py --gpus 1 --workload c5 --steps 10 --warmup 3 --no-cpu > gpurun_out/r5_bench_c5_n1.log 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655"
timeout 600 $TR bench.py --gpus 2 --workload c5 --steps 10 --warmup 3 > gpurun_out/r5_bench_c5_n2.log 2>&1
timeout 300 python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu > gpurun_out/r5_plain_c3.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 30 -c 40 --csv --log-file gpurun_out/r5_launches_c3.csv python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu > gpurun_out/r5_ncu_c3.log 2>&1
ls -la gpurun_out | tail -8
