#!/bin/bash
# GPU session 35 (8 GPUs): BASELINE configs[4] at its real size -- NLSE 3D cubic 1024^3 complex128, z-slabs over 8 B200
# (slab worker parity at 8 ranks first, then the C5 bench line at N=8)
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29671"
timeout 420 $TR tests/slab_worker.py > gpurun_out/r35_slab_worker_8gpu.log 2>&1; echo "slab worker rc=$?" >> gpurun_out/r35_slab_worker_8gpu.log
timeout 420 $TR bench.py --gpus 8 --workload c5 --steps 10 --warmup 3 > gpurun_out/r35_bench_c5_n8.log 2>&1; echo "bench rc=$?" >> gpurun_out/r35_bench_c5_n8.log
grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/r35_bench_*.log
tail -5 gpurun_out/r35_slab_worker_8gpu.log
