#!/bin/bash
# GPU session 3 (2 GPUs): full parity suite after the 2-pass Lanczos rewrite, slab worker, benches
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r3_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r3_pytest_gpu.log
for w in c3 c4 c1 c2_sv; do
  timeout 300 python bench.py --workload $w --steps 100 --warmup 5 --no-cpu > gpurun_out/r3_bench_$w.log 2>&1
done
timeout 300 python bench.py --steps 300 --warmup 10 --no-cpu > gpurun_out/r3_bench_c2_gautschi.log 2>&1
timeout 600 python bench.py --gpus 1 --workload c5 --steps 10 --warmup 3 --no-cpu > gpurun_out/r3_bench_c5_n1.log 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655"
timeout 600 $TR bench.py --gpus 2 --workload c5 --steps 10 --warmup 3 > gpurun_out/r3_bench_c5_n2.log 2>&1
timeout 300 python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu > gpurun_out/r3_plain_c3.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 30 -c 60 --csv --log-file gpurun_out/r3_launches_c3.csv python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu > gpurun_out/r3_ncu_c3.log 2>&1
ls -la gpurun_out | tail -15
