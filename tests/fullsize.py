"""Full-size parity cases (VERDICT r01 "next" item 1): every BASELINE config at the
grid size and launch shape the bench numbers come from, CUDA path vs the CPU oracle.

The oracle needs tens of seconds to minutes per case on one core, so all cases are started
together in a thread pool (ctypes releases the GIL; one core per case) the first time any of them
is asked for, and every test just waits for its own future.  TEST INFRASTRUCTURE ONLY.
"""
import concurrent.futures as cf
import os
import threading

import numpy as np

import ics
from nls_b200 import Desc, EQ, INT, LAP, FIELD

_lock = threading.Lock()
_pool = None
_futures = {}


slab_ic = ics.slab_ic


def c4_batch():
    """the C4 ensemble members checked against the oracle: trajectories t of the seed-10000+t family.
    One per-GPU batch of 64 (what the bench runs); positions 0, 17, 62, 63 hold t = 0, 17, 255, 511."""
    ts = list(range(62)) + [255, 511]
    return ts, (0, 17, 62, 63)


# name -> dict(desc kwargs, state, dt, checkpoints (step counts at which U [and V] are recorded))
def cases():
    cs = {}
    L2 = 10.0
    for n, steps, cps in ((1024, 200, (50, 200)), (256, 4999, (1000, 4999))):
        cs[f"c2_gautschi_{n}"] = dict(kw=dict(nx=n, ny=n, L=L2, equation=EQ.SGE, integrator=INT.GAUTSCHI, krylov_m=10, dt=1e-3),
                                      u0=lambda n=n: ics.kink_2d(n, L2)[None], real=True, dt=1e-3, cps=cps)
    for n, steps, cps in ((1024, 2000, (200, 2000)), (256, 4999, (1000, 4999))):
        cs[f"c2_sv_{n}"] = dict(kw=dict(nx=n, ny=n, L=L2, equation=EQ.SGE, integrator=INT.SV, krylov_m=10, dt=1e-3),
                                u0=lambda n=n: ics.kink_2d(n, L2)[None], real=True, dt=1e-3, cps=cps)
    for n, cps in ((64, (50, 200)), (128, (20,)), (256, (2,))):
        cs[f"c3_{n}"] = dict(kw=dict(nx=n, ny=n, nz=n, L=6.0, equation=EQ.NLSE_CQ, laplacian=LAP.ANISO, krylov_m=12,
                                     s1=1.0, s2=-0.1),
                             u0=lambda n=n: ics.gaussian_bumps_3d(n, 6.0, seed=99)[None], real=False, dt=5e-4, cps=cps,
                             cfield=True)
    ts, _ = c4_batch()
    cs["c4_512"] = dict(kw=dict(nx=512, ny=512, L=10.0, equation=EQ.NLSE_CUBIC, krylov_m=10),
                        u0=lambda: np.stack([ics.gaussian_bumps_2d(512, 10.0, seed=10000 + t) for t in ts]),
                        real=False, dt=1e-3, cps=(200,), oracle_members=c4_batch()[1])
    # C5: the slab workload's equation/IC; oracle-affordable cube and one slab-shaped grid (1024 x 1024 x 16:
    # the 16 MiB planes and 1024-wide rows of the real slabs, 256^3-many sites)
    cs["c5_128"] = dict(kw=dict(nx=128, ny=128, nz=128, L=10.0, equation=EQ.NLSE_CUBIC, krylov_m=10),
                        u0=lambda: slab_ic(128, 0, 128, 128, 10.0)[None], real=False, dt=2e-4, cps=(10,))
    cs["c5_1024x16"] = dict(kw=dict(nx=1024, ny=1024, nz=16, L=10.0, equation=EQ.NLSE_CUBIC, krylov_m=10),
                            u0=lambda: slab_ic(1024, 56, 16, 128, 10.0)[None], real=False, dt=2e-4, cps=(2,))
    return cs


def make_desc(case, batch, flags=0):
    kw = dict(case["kw"])
    nx, ny, nz = kw.pop("nx"), kw.pop("ny"), kw.pop("nz", 1)
    return Desc.make(nx, ny, nz, batch=batch, flags=flags, **kw)


def run(solver_cls, case, u0, flags=0, members=None):
    """advance `u0` (batch, ...) with solver_cls (Engine or Oracle); returns {step: (U[, V])}."""
    if members is not None:
        u0 = u0[list(members)]
    d = make_desc(case, u0.shape[0], flags)
    s = solver_cls(d)
    s.set(FIELD.M, np.ones(d.n))
    if case.get("cfield"):
        s.set(FIELD.C, np.ones(d.n))
    s.set(FIELD.U, u0)
    if case["real"]:
        s.set(FIELD.V, np.zeros_like(u0))
    out, done = {}, 0
    for cp in case["cps"]:
        s.step(case["dt"], cp - done)
        done = cp
        out[cp] = (s.get(FIELD.U), s.get(FIELD.V)) if case["real"] else (s.get(FIELD.U),)
    return out


def _oracle_job(name, case):
    import orc
    u0 = case["u0"]()
    return run(orc.Oracle, case, u0, members=case.get("oracle_members"))


def start_all():
    """start every case's oracle run (idempotent, returns immediately)."""
    global _pool
    with _lock:
        if _pool is not None:
            return
        import orc
        orc.build()
        orc.lib()
        ncpu = len(os.sched_getaffinity(0))
        orc.Oracle.set_threads(max(1, min(4, ncpu // 4)))  # only the C4 case has batch > 1
        _pool = cf.ThreadPoolExecutor(max_workers=max(2, min(12, ncpu)))
        allc = cases()
        for k in sorted(allc, key=lambda k: -_cost(allc[k])):  # longest first
            _futures[k] = _pool.submit(_oracle_job, k, allc[k])


def oracle_result(name):
    """result of the named case's oracle run (blocks until that case is done)."""
    start_all()
    return _futures[name].result()


def _cost(case):
    kw = case["kw"]
    n = kw["nx"] * kw["ny"] * kw.get("nz", 1)
    per = {INT.SV: 0.03, INT.GAUTSCHI: 0.8}.get(kw.get("integrator", INT.SS2), 0.65 if kw.get("nz", 1) == 1 else 1.1)
    return n * max(case["cps"]) * per
