#!/bin/bash
# GPU session 41 (2 GPUs): full GPU suite on the final library with the 2-GPU tests enabled (slab worker, --devices 2 CLI split)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r41_pytest_gpu_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r41_pytest_gpu_2gpu.log
tail -4 gpurun_out/r41_pytest_gpu_2gpu.log
