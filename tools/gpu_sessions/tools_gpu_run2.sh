#!/bin/bash
# GPU session 2 (2 GPUs): slab parity worker, c5 slab bench at N=2 and N=1, ensemble bench at N=2, c3 launch list with DRAM bytes
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2_gpus.txt
nvidia-smi topo -m >> gpurun_out/r2_gpus.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655"
timeout 600 $TR tests/slab_worker.py > gpurun_out/r2_slab_worker.log 2>&1; echo "rc=$?" >> gpurun_out/r2_slab_worker.log
timeout 600 $TR bench.py --gpus 2 --workload c5 --steps 10 --warmup 3 > gpurun_out/r2_bench_c5_n2.log 2>&1
timeout 600 python bench.py --gpus 1 --workload c5 --steps 10 --warmup 3 --no-cpu > gpurun_out/r2_bench_c5_n1.log 2>&1
timeout 600 $TR bench.py --gpus 2 --steps 200 --warmup 10 > gpurun_out/r2_bench_c2_n2.log 2>&1
timeout 300 python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu > gpurun_out/r2_plain_c3.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 40 -c 80 --csv --log-file gpurun_out/r2_launches_c3.csv python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu > gpurun_out/r2_ncu_c3.log 2>&1
ls -la gpurun_out
