#!/bin/bash
# GPU session 27 (4 GPUs): slab worker at 4 ranks, C5 N=4, default bench N=4
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29655"
timeout 600 $TR tests/slab_worker.py > gpurun_out/r27_slab_worker_4gpu.log 2>&1; echo "slab worker rc=$?" >> gpurun_out/r27_slab_worker_4gpu.log
timeout 600 $TR bench.py --gpus 4 --workload c5 --steps 10 --warmup 3 > gpurun_out/r27_bench_c5_n4.log 2>&1
timeout 600 $TR bench.py --gpus 4 > gpurun_out/r27_bench_c2_gautschi_n4.log 2>&1
grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/r27_bench_*.log
tail -5 gpurun_out/r27_slab_worker_4gpu.log
