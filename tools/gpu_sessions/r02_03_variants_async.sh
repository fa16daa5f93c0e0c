#!/bin/bash
# GPU session r02-03 (1 GPU): A/B timing of the persistent-kernel variants (A: relaxed polling + fence, register QL;
# B: acquire-load polling; C: A with the shared-memory QL; T: A with the phase trace), the asynchronous transfer
# tests, and the default bench line with the asynchronous e2e.
set -x
mkdir -p gpurun_out
for v in A B C; do
  for wl in c1 c2_gautschi; do
    NLS_B200_LIB=$PWD/nonlinear-solvers_b200/variants/libnls_$v.so timeout 300 python bench.py --workload $wl --no-others --no-cpu --steps 400 --warmup 20 > gpurun_out/r02c_var${v}_$wl.log 2>&1
    python - <<PY
import json
for ln in open("gpurun_out/r02c_var${v}_$wl.log"):
    if ln.startswith("{"):
        d = json.loads(ln); print("variant $v $wl ms/step", d["ms_per_step"], "e2e/value", d["e2e"]["value"]/d["value"], "blocking", d["e2e"]["blocking_value"]/d["value"])
PY
  done
done
NLS_B200_LIB=$PWD/nonlinear-solvers_b200/variants/libnls_T.so timeout 300 python bench.py --workload c2_gautschi --no-others --no-cpu --steps 10 --warmup 3 > gpurun_out/r02c_trace_c2_gautschi.log 2>&1
grep -c PKT gpurun_out/r02c_trace_c2_gautschi.log
( time timeout 1200 python -m pytest tests/test_gpu_async.py tests/test_gpu_parity.py -m gpu -q -x ) > gpurun_out/r02c_pytest_async.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02c_pytest_async.log
tail -15 gpurun_out/r02c_pytest_async.log
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 ) > gpurun_out/r02c_bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/r02c_bench.log
tail -4 gpurun_out/r02c_bench.log | cut -c1-3000
