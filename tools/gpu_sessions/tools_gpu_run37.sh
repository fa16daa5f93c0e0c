#!/bin/bash
# GPU session 37 (1 GPU): final library of the round -- full GPU suite, smoke, default bench, energy reduction timing after the 32-bit index change
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r37_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r37_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r37_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r37_smoke.log
timeout 600 python bench.py > gpurun_out/r37_bench_c2_gautschi.log 2>&1
timeout 300 python - > gpurun_out/r37_energy_reduce_time.log 2>&1 <<'PY'
import time, numpy as np
from nls_b200 import Desc, Engine, EQ, INT, FIELD, RED
d = Desc.make(256, 256, 256, L=10.0, equation=EQ.NLSE_CUBIC, krylov_m=4)
e = Engine(d)
rng = np.random.default_rng(0)
e.set(FIELD.U, rng.standard_normal(d.n) + 1j * rng.standard_normal(d.n))
for what, nm in [(RED.MASS, "mass"), (RED.ENERGY, "energy")]:
    for _ in range(3):
        e.reduce(what)
    t0 = time.perf_counter()
    for _ in range(20):
        r = e.reduce(what)
    dt = (time.perf_counter() - t0) / 20
    print(f"c3 256^3 complex {nm}: {dt*1e6:.1f} us per call incl. launch+D2H+sync, {d.n*16/dt/1e9:.0f} GB/s algorithmic, value {r[0]:.6e}")
PY
tail -3 gpurun_out/r37_pytest_gpu.log; tail -2 gpurun_out/r37_smoke.log; cat gpurun_out/r37_energy_reduce_time.log; grep -o '"ms_per_step": [0-9.]*' gpurun_out/r37_bench_c2_gautschi.log
