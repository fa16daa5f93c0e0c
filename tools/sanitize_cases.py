"""Small invocations of every kernel family, for compute-sanitizer (memcheck / racecheck / synccheck):
the persistent cooperative kernels (grid barrier, partial-sum gather, in-kernel tridiagonal solve), the
multi-kernel graph path (ticket reduction), Stoermer-Verlet, sEWI, the resampling / energy kernels.
Usage: compute-sanitizer --tool <tool> python tools/sanitize_cases.py [case ...]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import nls_b200  # noqa: E402
from nls_b200 import Desc, Engine, EQ, INT, LAP, FIELD, RED, FN  # noqa: E402
import ics  # noqa: E402

NOPK = nls_b200.FLAG_NO_PERSISTENT | nls_b200.FLAG_NO_GRAPH


def run(name, d, u0, steps, dt=1e-3, v0=None):
    e = Engine(d)
    e.set(FIELD.M, np.ones(d.n))
    if d.laplacian == LAP.ANISO:
        e.set(FIELD.C, np.ones(d.n))
    e.set(FIELD.U, u0)
    if v0 is not None:
        e.set(FIELD.V, v0)
    e.step(dt, steps)
    u = e.get()
    m = e.reduce(RED.MASS)[0]
    en = e.reduce(RED.ENERGY)[0]
    r = e.get_resampled(16, 16, 16 if d.nz > 1 else 1)
    assert np.isfinite(u).all() and np.isfinite(m) and np.isfinite(en) and np.isfinite(r).all()
    print(f"{name}: ok ({e.launch_count()} launches, mass {m:.6e})", flush=True)
    e.close()


def main():
    want = set(sys.argv[1:])
    n = 48
    u2 = ics.gaussian_bumps_2d(n, 10.0)[None]
    k2 = ics.kink_2d(n, 10.0)[None]
    u3 = ics.gaussian_bumps_3d(20, 6.0)[None]
    cases = {
        "pk_nlse": lambda: run("pk_nlse", Desc.make(n, n, L=10.0, equation=EQ.NLSE_CUBIC, krylov_m=10), u2, 3),
        "pk_nlse_cq3d": lambda: run("pk_nlse_cq3d", Desc.make(20, 20, 20, L=6.0, equation=EQ.NLSE_CQ, laplacian=LAP.ANISO,
                                                              krylov_m=12, s1=1.0, s2=-0.1), u3, 2, 5e-4),
        "pk_gautschi": lambda: run("pk_gautschi", Desc.make(n, n, L=10.0, equation=EQ.SGE, integrator=INT.GAUTSCHI,
                                                            krylov_m=10, dt=1e-3), k2, 3, v0=np.zeros_like(k2)),
        "pk_sv": lambda: run("pk_sv", Desc.make(n, n, L=10.0, equation=EQ.SGE, integrator=INT.SV, dt=1e-3,
                                                flags=nls_b200.FLAG_PK_SV), k2, 10, v0=np.zeros_like(k2)),
        "mk_nlse": lambda: run("mk_nlse", Desc.make(n, n, L=10.0, equation=EQ.NLSE_CUBIC, krylov_m=10, flags=NOPK), u2, 3),
        "mk_nlse_batch": lambda: run("mk_nlse_batch", Desc.make(n, n, L=10.0, equation=EQ.NLSE_SAT, krylov_m=8, kappa=0.3,
                                                                batch=3, flags=NOPK), np.repeat(u2, 3, axis=0), 2),
        "mk_cq3d": lambda: run("mk_cq3d", Desc.make(20, 20, 20, L=6.0, equation=EQ.NLSE_CQ, laplacian=LAP.ANISO, krylov_m=12,
                                                    s1=1.0, s2=-0.1, flags=NOPK), u3, 2, 5e-4),
        "mk_gautschi": lambda: run("mk_gautschi", Desc.make(n, n, L=10.0, equation=EQ.KGE, integrator=INT.GAUTSCHI, krylov_m=10,
                                                            dt=1e-3, op_sign=-1.0, flags=NOPK), 0.2 * k2, 2, v0=np.zeros_like(k2)),
        "mk_sv": lambda: run("mk_sv", Desc.make(n, n, L=10.0, equation=EQ.SGE, integrator=INT.SV, dt=1e-3, flags=NOPK), k2, 10,
                             v0=np.zeros_like(k2)),
        "mk_sewi": lambda: run("mk_sewi", Desc.make(n, n, L=10.0, equation=EQ.NLSE_CUBIC, integrator=INT.SEWI, krylov_m=8,
                                                    flags=NOPK), u2, 3),
    }
    for k, fn in cases.items():
        if not want or k in want:
            fn()
    print("SANITIZE_CASES_OK", flush=True)


if __name__ == "__main__":
    main()
