#!/bin/bash
# GPU session 33 (1 GPU): NLS_RED_ENERGY (fused drift diagnostics) + --diag driver flag; full GPU suite on the new library;
# wall time of the energy reduction at C3 (256^3 complex) and C2 (1024^2 real) sizes
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r33_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r33_pytest_gpu.log
timeout 300 python - > gpurun_out/r33_energy_reduce_time.log 2>&1 <<'PY'
import time, numpy as np, torch
from nls_b200 import Desc, Engine, EQ, INT, FIELD, RED
for name, d, bytes_per_site in [
    ("c3 256^3 complex", Desc.make(256, 256, 256, L=10.0, equation=EQ.NLSE_CUBIC, krylov_m=4), 16),
    ("c2 1024^2 real", Desc.make(1024, 1024, L=10.0, equation=EQ.SGE, integrator=INT.GAUTSCHI, dt=1e-3), 16)]:
    e = Engine(d)
    rng = np.random.default_rng(0)
    u = rng.standard_normal(d.n) if bytes_per_site == 16 and d.equation == EQ.SGE else rng.standard_normal(d.n) + 1j * rng.standard_normal(d.n)
    e.set(FIELD.U, u)
    if d.equation == EQ.SGE:
        e.set(FIELD.V, rng.standard_normal(d.n))
    for what, nm in [(RED.MASS, "mass"), (RED.ENERGY, "energy")]:
        for _ in range(3):
            e.reduce(what)
        t0 = time.perf_counter()
        for _ in range(20):
            r = e.reduce(what)
        dt = (time.perf_counter() - t0) / 20
        b = d.n * (16 if d.equation != EQ.SGE else (16 if what == RED.ENERGY else 8))
        print(f"{name} {nm}: {dt*1e6:.1f} us per call incl. launch+D2H+sync, {b/dt/1e9:.0f} GB/s algorithmic, value {r[0]:.6e}")
PY
tail -3 gpurun_out/r33_pytest_gpu.log; cat gpurun_out/r33_energy_reduce_time.log
