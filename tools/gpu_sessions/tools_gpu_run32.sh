#!/bin/bash
# GPU session 32 (2 GPUs): ensemble-file drivers -- CLI parity tests incl. the --devices split,
# and the C4-shaped ensemble through the CLI (64 x 512^2, 100 steps) on 1 and 2 devices
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_drivers.py -m gpu -q > gpurun_out/r32_pytest_drivers.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r32_pytest_drivers.log
python - <<'PY'
import numpy as np
rng = np.random.default_rng(3)
B, n, L = 64, 512, 10.0
x = np.linspace(-L, L, n); X, Y = np.meshgrid(x, x)
u = np.empty((B, n, n), dtype=np.complex128)
for b in range(B):
    cx, cy, kx, ky = rng.uniform(-3, 3, 4)
    u[b] = np.exp(-((X - cx) ** 2 + (Y - cy) ** 2)) * np.exp(1j * (kx * X + ky * Y))
np.save("/tmp/ens_u0.npy", u)
PY
D=nonlinear-solvers_b200/host/bin
for G in 1 2; do
  NLS_VERBOSE=1 timeout 600 $D/nlse_ensemble_driver 512 512 10.0 10.0 /tmp/ens_u0.npy /tmp/ens_t$G.npy 0.1 101 4 --devices $G > gpurun_out/r32_ensemble_cli_g$G.log 2>&1; echo "rc=$?" >> gpurun_out/r32_ensemble_cli_g$G.log
done
python -c "
import numpy as np
a=np.load('/tmp/ens_t1.npy'); b=np.load('/tmp/ens_t2.npy')
print('shape',a.shape,'G=2 vs G=1 rel-L2', np.linalg.norm(a-b)/np.linalg.norm(a))" > gpurun_out/r32_ensemble_cli_cmp.log 2>&1
cat gpurun_out/r32_*.log | tail -30
