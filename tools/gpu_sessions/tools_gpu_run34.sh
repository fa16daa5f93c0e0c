#!/bin/bash
# GPU session 34 (1 GPU): nls_get_field_resampled (on-device downsampling) + --downsample driver flag;
# full GPU suite on the new library; snapshot egress time full vs downsampled (C2 1024^2 -> 128^2, C3 256^3 -> 64^3)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r34_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r34_pytest_gpu.log
timeout 300 python - > gpurun_out/r34_egress_time.log 2>&1 <<'PY'
import time, numpy as np
from nls_b200 import Desc, Engine, EQ, INT, FIELD
rng = np.random.default_rng(0)
for name, d, tgt in [
    ("c2 1024^2 real -> 128^2", Desc.make(1024, 1024, L=10.0, equation=EQ.SGE, integrator=INT.GAUTSCHI, dt=1e-3), (128, 128, 1)),
    ("c1-like 512^2 complex x64 -> 128^2", Desc.make(512, 512, L=10.0, equation=EQ.NLSE_CUBIC, batch=64), (128, 128, 1)),
    ("c3 256^3 complex -> 64^3", Desc.make(256, 256, 256, L=10.0, equation=EQ.NLSE_CUBIC, krylov_m=4), (64, 64, 64))]:
    e = Engine(d)
    u = rng.standard_normal((d.batch, d.n))
    if d.equation != EQ.SGE:
        u = u + 1j * rng.standard_normal((d.batch, d.n))
    e.set(FIELD.U, u)
    out_full = np.empty(e._state_shape(), dtype=d.dtype)
    def t(f, reps=10):
        f(); f()
        t0 = time.perf_counter()
        for _ in range(reps):
            f()
        return (time.perf_counter() - t0) / reps
    tf = t(lambda: e.get_into(FIELD.U, out_full))
    tr = t(lambda: e.get_resampled(*tgt))
    small = e.get_resampled(*tgt)
    print(f"{name}: full snapshot D2H {tf*1e3:.3f} ms ({out_full.nbytes/2**20:.1f} MiB, pageable host), "
          f"resampled on device {tr*1e3:.3f} ms ({small.nbytes/2**20:.2f} MiB)  -> {tf/tr:.1f}x")
PY
timeout 300 python bench.py --no-cpu > gpurun_out/r34_bench_c2_gautschi.log 2>&1
tail -3 gpurun_out/r34_pytest_gpu.log; cat gpurun_out/r34_egress_time.log; grep -o '"ms_per_step": [0-9.]*' gpurun_out/r34_bench_c2_gautschi.log
