"""Generates the committed golden vectors in this directory.

The reference ships no golden vectors and its CPU path cannot be built here (Eigen3 absent), so the
goldens are outputs of the CPU oracle (oracle/nls_oracle.cpp, a restatement of the reference's
host algorithm) AFTER it has been pinned by tests/test_oracle.py against scipy's expm_multiply (the
reference's own external check, nlsolvers/scipy-test/check_krylov_compute.py:41-49), dense eigh
evaluation and the independent numpy twin.  One extra fixture is scipy's own answer for the
check_krylov_compute set-up, so the anchor itself travels to the GPU box.

    python tests/golden/make_golden.py          # rewrites tests/golden/*.npz
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import nls_b200  # noqa: E402
from nls_b200 import Desc, EQ, INT, LAP, FN, FIELD  # noqa: E402
import ics  # noqa: E402

# name -> (grid, L, Desc kwargs, steps, dt, kind of initial data)
CASES = {
    "nlse_cubic_2d": ((48, 48, 1), 10.0, dict(equation=EQ.NLSE_CUBIC, krylov_m=10), 25, 1e-3),
    "nlse_cq_2d": ((40, 40, 1), 10.0, dict(equation=EQ.NLSE_CQ, krylov_m=10, s1=1.0, s2=-0.1), 15, 1e-3),
    "nlse_sat_2d": ((40, 40, 1), 10.0, dict(equation=EQ.NLSE_SAT, krylov_m=10, kappa=0.3), 15, 1e-3),
    "nlse_cq_3d_aniso": ((16, 16, 16), 6.0, dict(equation=EQ.NLSE_CQ, laplacian=LAP.ANISO, krylov_m=12, s1=1.0, s2=-0.1),
                         8, 5e-4),
    "nlse_sewi_2d_host": ((40, 40, 1), 10.0, dict(equation=EQ.NLSE_CUBIC, integrator=INT.SEWI, krylov_m=10,
                                                 flags=nls_b200.FLAG_SEWI_HOST2D), 10, 1e-3),
    "nlse_sewi_3d": ((14, 14, 14), 6.0, dict(equation=EQ.NLSE_CUBIC, integrator=INT.SEWI, laplacian=LAP.ANISO, krylov_m=10),
                     8, 1e-3),
    "sge_gautschi_2d": ((48, 48, 1), 10.0, dict(equation=EQ.SGE, integrator=INT.GAUTSCHI, krylov_m=10), 12, 1e-3),
    "sge_sv_2d": ((48, 48, 1), 10.0, dict(equation=EQ.SGE, integrator=INT.SV), 100, 1e-3),
    "kge_gautschi_2d": ((40, 40, 1), 10.0, dict(equation=EQ.KGE, integrator=INT.GAUTSCHI, krylov_m=10, op_sign=-1.0), 12, 1e-3),
    "kge_sv_3d_aniso": ((14, 14, 14), 4.0, dict(equation=EQ.KGE, integrator=INT.SV, laplacian=LAP.ANISO), 30, 1e-3),
}


def inputs(name):
    (nx, ny, nz), L, kw, steps, dt = CASES[name]
    d = Desc.make(nx, ny, nz, L=L, dt=dt, **kw)
    rng = np.random.default_rng(2024)
    mf = 1.0 + 0.2 * rng.standard_normal(d.n)
    cf = 1.0 + 0.1 * rng.standard_normal(d.n)
    if d.is_complex:
        u0 = ics.gaussian_bumps_3d(nx, L) if nz > 1 else ics.gaussian_bumps_2d(nx, L)
        v0 = None
    else:
        u0 = 0.3 * np.abs(ics.gaussian_bumps_3d(nx, L)) if nz > 1 else ics.kink_2d(nx, L)
        if kw["equation"] == EQ.KGE and nz == 1:
            u0 = 0.2 * u0
        v0 = 0.05 * np.cos(u0)
    return d, u0, v0, mf, cf, steps, dt


def run(solver, name):
    """drive any object with the Engine/Oracle surface through a golden case"""
    d, u0, v0, mf, cf, steps, dt = inputs(name)
    s = solver(d)
    if d.laplacian == LAP.ANISO:
        s.set(FIELD.C, cf)
    s.set(FIELD.M, mf)
    s.set(FIELD.U, u0)
    if v0 is not None:
        s.set(FIELD.V, v0)
    s.step(dt, steps)
    out = {"u": s.get(FIELD.U)[0]}
    if not d.is_complex:
        out["v"] = s.get(FIELD.V)[0]
    return out


def matfunc_cases():
    """f(tL)x on white noise (device/test_matfunc_*.cpp set-up: mt19937-style N(0,1) inputs, t = 1e-2)"""
    out = []
    for fn, t, eq, integ, sign in [(FN.EXP, 1e-2j, EQ.NLSE_CUBIC, INT.SS2, 1.0), (FN.SINC, 1e-2, EQ.NLSE_CUBIC, INT.SS2, 1.0),
                                   (FN.COS_SQRT, 1e-2, EQ.SGE, INT.GAUTSCHI, -1.0), (FN.SINC2_HALF, 1e-2, EQ.SGE, INT.GAUTSCHI, -1.0),
                                   (FN.ID_SQRT, 1e-2, EQ.SGE, INT.GAUTSCHI, -1.0), (FN.MOD_COSINE, 1e-2, EQ.SGE, INT.GAUTSCHI, -1.0),
                                   (FN.SINC2_SQRT, 1e-2, EQ.SGE, INT.GAUTSCHI, -1.0)]:
        d = Desc.make(36, 36, L=5.0, equation=eq, integrator=integ, krylov_m=10, op_sign=sign)
        rng = np.random.default_rng(42)
        x = rng.standard_normal((1, d.n))
        if d.is_complex:
            x = x + 1j * rng.standard_normal((1, d.n))
        out.append((f"matfunc_fn{fn}", d, fn, t, x))
    return out


def main():
    import orc
    from scipy.sparse.linalg import expm_multiply
    for name in CASES:
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **run(orc.Oracle, name))
    mf = {}
    for key, d, fn, t, x in matfunc_cases():
        mf[key] = orc.Oracle(d).matfunc(fn, t, x)[0]
    np.savez_compressed(os.path.join(HERE, "matfuncs.npz"), **mf)
    # scipy's own answer for the reference's external check (check_krylov_compute.py:41-49)
    n = 16
    d = Desc.make(n, n, n, L=2.0, laplacian=LAP.ANISO, krylov_m=30)
    o = orc.Oracle(d)
    o.set(FIELD.C, np.ones(d.n))
    u0 = ics.gaussian_bumps_3d(n, 2.0, K=1).ravel()
    ref = expm_multiply(1e-2j * o.csr().astype(complex), u0)
    np.savez_compressed(os.path.join(HERE, "scipy_expm_3d.npz"), u0=u0, expm=ref)
    print("wrote", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))


if __name__ == "__main__":
    main()
