#!/bin/bash
# GPU session 24 (1 GPU): ncu --set full capture of the final persistent Gautschi kernel
set -x
mkdir -p gpurun_out
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/r24_plain.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_pk_gautschi -s 2 -c 1 -f -o gpurun_out/r24_pk_gautschi_full python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/r24_ncu.log 2>&1
ls -la gpurun_out | grep r24
