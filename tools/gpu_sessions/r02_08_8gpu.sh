#!/bin/bash
# GPU session r02-08 (8 GPUs, one box): slab parity worker at 8 ranks, the default bench (C5 weak-scaling series) at
# N = 1, 2, 4, 8 the way the driver launches it, and the strong-scaling series of the C3 grid (256^3 cut into 1/2/4/8 slabs).
set -x
mkdir -p gpurun_out
nvidia-smi -L | wc -l
run() { # N workload tag extra...
  local n=$1 wl=$2 tag=$3; shift 3
  if [ $n = 1 ]; then
    timeout 200 python bench.py --gpus 1 --steps 10 --warmup 3 --workload $wl "$@" > gpurun_out/r02h_${tag}_n$n.log 2>&1
  else
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) bench.py --gpus $n --steps 10 --warmup 3 --workload $wl "$@" > gpurun_out/r02h_${tag}_n$n.log 2>&1
  fi
  echo "rc=$?" >> gpurun_out/r02h_${tag}_n$n.log
}
run 8 c5 bench_c5 --no-cpu
for n in 4 2 1; do run $n c5 bench_c5 --no-cpu --no-others; done  # box time is charged 8x: the others block only at N = 8
for n in 1 2 4 8; do run $n c3_strong bench_c3_strong --no-cpu --no-others; done
python - <<'PY'
import json, glob
for tag in ("bench_c5", "bench_c3_strong"):
    base = None
    for n in (1, 2, 4, 8):
        for ln in open(f"gpurun_out/r02h_{tag}_n{n}.log"):
            if ln.startswith("{"):
                d = json.loads(ln)
                if n == 1: base = d
                eff = (base["ms_per_step"] / d["ms_per_step"]) if tag == "bench_c5" else (base["ms_per_step"] / d["ms_per_step"] / n)
                print(tag, "N", n, "ms/step %.3f" % d["ms_per_step"], "value %.4g" % d["value"], "e2e %.4g" % d["e2e"]["value"], "eff %.3f" % eff,
                      d["clocks"]["reasons"], {k: round(v["ms_per_step"], 3) for k, v in d.get("others", {}).items()})
PY
# last (box time is charged 8x): the slab parity worker at 8 ranks on the final kernels
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29650 tests/slab_worker.py > gpurun_out/r02h_slab_worker_8gpu.log 2>&1; echo "rc=$?" >> gpurun_out/r02h_slab_worker_8gpu.log
tail -4 gpurun_out/r02h_slab_worker_8gpu.log
