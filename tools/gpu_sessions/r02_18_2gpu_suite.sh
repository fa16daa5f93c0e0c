#!/bin/bash
# GPU session r02-18 (2 GPUs): the whole GPU suite on the final library with the 2-GPU cases not skipped (slab worker at
# 2 ranks incl. the 256^3 slab-count invariance with the tensor-map pass, ensemble CLI), then the C5 line at N=2.
set -x
mkdir -p gpurun_out
nvidia-smi -L | wc -l
( timeout 1500 python -m pytest tests -m gpu -q -x ) > gpurun_out/r02r_pytest_gpu_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02r_pytest_gpu_2gpu.log
tail -4 gpurun_out/r02r_pytest_gpu_2gpu.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --no-cpu --no-others > gpurun_out/r02r_bench_c5_n2.log 2>&1; echo "rc=$?" >> gpurun_out/r02r_bench_c5_n2.log
grep '^{' gpurun_out/r02r_bench_c5_n2.log | cut -c1-500
