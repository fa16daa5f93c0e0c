#!/bin/bash
# GPU session 26 (2 GPUs): full GPU suite incl. the slab worker, C5 N=1/N=2, default bench N=1/N=2
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r26_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r26_pytest_gpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655"
timeout 600 python bench.py --gpus 1 --workload c5 --steps 10 --warmup 3 --no-cpu > gpurun_out/r26_bench_c5_n1.log 2>&1
timeout 600 $TR bench.py --gpus 2 --workload c5 --steps 10 --warmup 3 > gpurun_out/r26_bench_c5_n2.log 2>&1
timeout 600 python bench.py --gpus 1 --no-cpu > gpurun_out/r26_bench_c2_gautschi_n1.log 2>&1
timeout 600 $TR bench.py --gpus 2 > gpurun_out/r26_bench_c2_gautschi_n2.log 2>&1
timeout 600 $TR bench.py --gpus 2 --impl reference --steps 2 --warmup 1 > gpurun_out/r26_bench_reference_n2.log 2>&1
grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/r26_bench_*.log
