#!/bin/bash
# GPU session r02-06 (2 GPUs): whole GPU suite with the 2-GPU tests enabled (z-slab worker incl. the 256^3 slab-count
# invariance cases, --devices 2 CLI split), then the default bench (C5 weak-scaling series) at N = 1 and N = 2.
set -x
mkdir -p gpurun_out
nvidia-smi -L
( time timeout 2700 python -m pytest tests -m gpu -q -x -rA --durations=10 ) > gpurun_out/r02f_pytest_gpu_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02f_pytest_gpu_2gpu.log
grep -E "passed|failed|rc=|SLAB|slab x2" gpurun_out/r02f_pytest_gpu_2gpu.log | tail -40
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 --no-others ) > gpurun_out/r02f_bench_n1.log 2>&1; echo "rc=$?" >> gpurun_out/r02f_bench_n1.log
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 ) > gpurun_out/r02f_bench_n2.log 2>&1; echo "rc=$?" >> gpurun_out/r02f_bench_n2.log
python - <<'PY'
import json
for n in (1, 2):
    for ln in open(f"gpurun_out/r02f_bench_n{n}.log"):
        if ln.startswith("{"):
            d = json.loads(ln)
            print("N", n, "ms/step", d["ms_per_step"], "value", d["value"], "e2e", d["e2e"]["value"], "blocking", d["e2e"]["blocking_value"], d["config"]["parallelism"], {k: (v["ms_per_step"], v["value"]) for k, v in d["others"].items()})
PY
tail -3 gpurun_out/r02f_bench_n2.log | cut -c1-600
