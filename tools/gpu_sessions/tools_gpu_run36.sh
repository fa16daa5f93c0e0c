#!/bin/bash
# GPU session 36 (8 GPUs): slab worker parity at 8 ranks (world-aware plane counts + the thinnest legal slabs: 2 planes per rank)
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29672"
timeout 420 $TR tests/slab_worker.py > gpurun_out/r36_slab_worker_8gpu.log 2>&1; echo "slab worker rc=$?" >> gpurun_out/r36_slab_worker_8gpu.log
grep "slab x8\|SLAB_\|rc=" gpurun_out/r36_slab_worker_8gpu.log | tail -50
