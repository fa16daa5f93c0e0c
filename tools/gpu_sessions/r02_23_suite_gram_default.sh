#!/bin/bash
# GPU session r02-23 (1 GPU): the whole GPU suite with the Gram-delayed form as the multi-kernel default.
mkdir -p gpurun_out
( timeout 215 python -m pytest tests -m gpu -q -x ) > gpurun_out/r02w_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02w_pytest_gpu.log
tail -6 gpurun_out/r02w_pytest_gpu.log
