"""Phase-timing trace of the persistent Gautschi kernel (debug build with -DPK_TRACE, see nls_persist.cuh).
Usage: NLS_B200_LIB=nonlinear-solvers_b200/variants/libnls_trace.so python tools/pk_trace.py > trace.txt"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import nls_b200
from nls_b200 import Desc, Engine, EQ, INT, FIELD
import ics
n, L, dt = 1024, 10.0, 1e-3
d = Desc.make(n, n, L=L, equation=EQ.SGE, integrator=INT.GAUTSCHI, krylov_m=10, dt=dt)
u0 = ics.kink_2d(n, L)
e = Engine(d)
e.set(FIELD.U, u0); e.set(FIELD.V, np.zeros_like(u0)); e.set(FIELD.M, np.ones(d.n))
e.step(dt, 20)
e.step(dt, 5)
print("done", float(np.abs(e.get()).max()), flush=True)
