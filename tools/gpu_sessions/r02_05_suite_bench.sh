#!/bin/bash
# GPU session r02-05 (1 GPU): whole GPU suite on the library with the counter barrier restored (register QL kept,
# temporally blocked SV, async transfers, B2 layer), then the default bench line.
set -x
mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -m gpu -q -x -rA --durations=10 ) > gpurun_out/r02e_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02e_pytest_gpu.log
grep -E "^(PASSED|FAILED|ERROR)|passed|failed|rc=" gpurun_out/r02e_pytest_gpu.log | grep -v "^PASSED" | tail -8
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 ) > gpurun_out/r02e_bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/r02e_bench.log
python - <<'PY'
import json
for ln in open("gpurun_out/r02e_bench.log"):
    if ln.startswith("{"):
        d = json.loads(ln)
        print("c5", d["ms_per_step"], d["value"], "e2e", d["e2e"]["value"], "blocking", d["e2e"]["blocking_value"], "frac", d["roofline"]["frac"])
        for k, v in d["others"].items():
            print(k, v["ms_per_step"], v["value"], "e2e", v["e2e"], "frac", v["roofline"]["frac"])
PY
