#!/bin/bash
# GPU session r02-02 (1 GPU): register-resident QL, fused barrier + all-reduce (sentinel polling), temporally
# blocked Stoermer-Verlet: parity (whole GPU suite) + timings of C1 / C2 Gautschi / C2 SV / sEWI-sized cases.
set -x
mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -m gpu -q -x -rA --durations=10 ) > gpurun_out/r02b_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02b_pytest_gpu.log
grep -E "passed|failed|error|rc=" gpurun_out/r02b_pytest_gpu.log | tail -8
for wl in c1 c2_gautschi c2_sv; do
  timeout 600 python bench.py --workload $wl --no-others --no-cpu --steps 400 --warmup 20 > gpurun_out/r02b_bench_$wl.log 2>&1; echo "rc=$?" >> gpurun_out/r02b_bench_$wl.log
  python - <<PY
import json
for ln in open("gpurun_out/r02b_bench_$wl.log"):
    if ln.startswith("{"):
        d = json.loads(ln); print("$wl", d["ms_per_step"], d["value"], d["e2e"]["value"], d["roofline"]["frac"])
PY
done
