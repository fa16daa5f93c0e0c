#!/bin/bash
# GPU session 38 (8 GPUs): BASELINE configs[3] at its real size -- 512 NLSE 2D 512^2 trajectories as 64 per GPU over 8 B200;
# and the default bench line (C2 Gautschi replicas) at N=8
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29673"
timeout 300 $TR bench.py --gpus 8 --workload c4 --no-cpu > gpurun_out/r38_bench_c4_n8.log 2>&1; echo "rc=$?" >> gpurun_out/r38_bench_c4_n8.log
timeout 300 $TR bench.py --gpus 8 --no-cpu > gpurun_out/r38_bench_c2_gautschi_n8.log 2>&1; echo "rc=$?" >> gpurun_out/r38_bench_c2_gautschi_n8.log
grep -H -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' gpurun_out/r38_bench_*.log
