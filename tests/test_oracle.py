"""CPU tests: pin the oracle (C++ restatement) against scipy's expm_multiply (the reference's
own external check, nlsolvers/scipy-test/check_krylov_compute.py:41-49), against dense
eigh evaluation of f(tL)u, and against the independent numpy twin."""
import numpy as np
import pytest
import scipy.sparse as sp
from scipy.sparse.linalg import expm_multiply

from conftest import rel_l2
import ics
import oracle_np as onp
from nls_b200 import Desc, EQ, INT, LAP, FN, FIELD


def test_laplacian_rows_2d(oracle_lib):
    d = Desc.make(16, 16, L=5.0)
    A = oracle_lib.Oracle(d).csr()
    rs = (A.sum(axis=1).A1 * d.dx * d.dy).reshape(16, 16)
    assert np.allclose(rs[1:-1, :], 0) and np.allclose(rs[0, 1:-1], 0)
    assert np.allclose(rs[[0, 0, -1, -1], [0, -1, 0, -1]], -1.0)  # corner quirk Q1
    assert abs(A - onp.lap2d_uniform(16, 16, d.dx, d.dy)).max() == 0
    assert abs(A - A.T).max() == 0


def test_laplacian_aniso_and_3d(oracle_lib):
    rng = np.random.default_rng(3)
    d = Desc.make(12, 12, L=3.0, laplacian=LAP.ANISO)
    c = rng.uniform(0.5, 1.5, d.n)
    o = oracle_lib.Oracle(d)
    o.set(FIELD.C, c)
    assert abs(o.csr() - onp.lap2d_aniso(12, 12, d.dx, d.dy, c)).max() < 1e-13
    d3 = Desc.make(7, 7, 7, L=2.0)
    A3 = oracle_lib.Oracle(d3).csr()
    assert abs(A3 - onp.lap3d(7, 7, 7, d3.dx)).max() == 0
    # wrap quirk Q2: (i, ny-1, k) <-> (i, 0, k+1)
    a, b = (0 * 7 + 6) * 7 + 3, (1 * 7 + 0) * 7 + 3
    assert A3[a, b] * d3.dx ** 2 == pytest.approx(1.0)
    d3a = Desc.make(6, 6, 6, L=2.0, laplacian=LAP.ANISO)
    c3 = rng.uniform(0.5, 1.5, d3a.n)
    o3 = oracle_lib.Oracle(d3a)
    o3.set(FIELD.C, c3)
    assert abs(o3.csr() - onp.lap3d(6, 6, 6, d3a.dx, c3)).max() < 1e-12


@pytest.mark.parametrize("m", [10, 20, 30])
def test_expm_vs_scipy_3d(oracle_lib, m):
    """test_scipy_matfunc.cpp:18-110 set-up: anisotropic 3D Laplacian, c=1, Gaussian IC, t=1e-2."""
    n = 20
    d = Desc.make(n, n, n, L=2.0, laplacian=LAP.ANISO, krylov_m=m)
    o = oracle_lib.Oracle(d)
    o.set(FIELD.C, np.ones(d.n))
    A = o.csr()
    u0 = ics.gaussian_bumps_3d(n, 2.0, K=1).ravel()
    got = o.matfunc(FN.EXP, 1j * 1e-2, u0)
    ref = expm_multiply(1j * 1e-2 * A.astype(complex), u0)
    assert rel_l2(got, ref) < 1e-8


def test_real_matfuncs_vs_dense_eigh(oracle_lib):
    n = 14
    d = Desc.make(n, n, L=3.0, equation=EQ.SGE, integrator=INT.GAUTSCHI, krylov_m=40)
    o = oracle_lib.Oracle(d)
    A = o.csr().toarray()
    lam, Q = np.linalg.eigh(A)
    u = ics.kink_2d(n, 3.0).ravel()
    t = 1e-2
    for fn, name in [(FN.COS_SQRT, "cos_sqrt"), (FN.SINC2_SQRT, "sinc2_sqrt"), (FN.ID_SQRT, "id_sqrt"),
                     (FN.SINC2_HALF, "sinc2_half"), (FN.MOD_COSINE, "mod_cosine")]:
        ref = Q @ (onp.f_table(name, t, lam) * (Q.T @ u))
        got = o.matfunc(fn, t, u)
        # id_sqrt is not smooth at 0: Krylov converges slowly, only a loose bound holds
        assert rel_l2(got, ref) < (5e-2 if fn == FN.ID_SQRT else 1e-9), name


@pytest.mark.parametrize("eq,kind", [(EQ.NLSE_CUBIC, "cubic"), (EQ.NLSE_CQ, "cq"), (EQ.NLSE_SAT, "sat")])
def test_nlse_steps_vs_numpy_twin(oracle_lib, eq, kind):
    n, L, dt = 24, 5.0, 1e-3
    d = Desc.make(n, n, L=L, equation=eq, krylov_m=10, s1=1.0, s2=-0.1, kappa=0.3)
    o = oracle_lib.Oracle(d)
    u0 = ics.gaussian_bumps_2d(n, L)
    mf = np.ones(d.n)
    o.set(FIELD.U, u0)
    o.set(FIELD.M, mf)
    o.step(dt, 5)
    A = onp.lap2d_uniform(n, n, d.dx, d.dy).astype(complex)
    u = u0.ravel().copy()
    for _ in range(5):
        u = nl = onp.nlse_step(A, u, mf, dt, 10, kind, 1.0, -0.1, 0.3)
        u = onp.neumann_bc(nl.reshape(n, n)).ravel()
    assert rel_l2(o.get()[0], u.reshape(n, n)) < 1e-12


@pytest.mark.parametrize("eq,kind,integ", [(EQ.SGE, "sge", INT.GAUTSCHI), (EQ.KGE, "kge", INT.GAUTSCHI),
                                           (EQ.PHI4, "phi4", INT.GAUTSCHI), (EQ.SGE, "sge", INT.SV),
                                           (EQ.SGE_SINH, "sge_sinh", INT.SV)])
def test_real_steps_vs_numpy_twin(oracle_lib, eq, kind, integ):
    n, L, dt = 24, 5.0, 1e-3
    d = Desc.make(n, n, L=L, equation=eq, integrator=integ, krylov_m=10, dt=dt)
    o = oracle_lib.Oracle(d)
    u0 = ics.kink_2d(n, L)
    v0 = 0.1 * np.cos(u0)
    mf = np.ones(d.n)
    o.set(FIELD.U, u0)
    o.set(FIELD.V, v0)
    o.set(FIELD.M, mf)
    o.step(dt, 4)
    A = onp.lap2d_uniform(n, n, d.dx, d.dy)
    u, up = u0.ravel().copy(), (u0 - dt * v0).ravel()
    for _ in range(4):
        if integ == INT.SV:
            u, up = onp.sv_step(A, u, up, mf, dt, kind)
        else:
            u, up = onp.gautschi_step(A, u, up, mf, dt, 10, kind)
        u = onp.neumann_bc(u.reshape(n, n)).ravel()
    assert rel_l2(o.get()[0], u.reshape(n, n)) < 1e-12
    assert rel_l2(o.get(FIELD.V)[0], ((u - up) / dt).reshape(n, n)) < 1e-8


def test_lanczos_last_diag_zero_and_orthogonality(oracle_lib):
    d = Desc.make(50, 50, L=5.0, equation=EQ.SGE, integrator=INT.GAUTSCHI, krylov_m=10)
    rng = np.random.default_rng(42)
    x = rng.standard_normal(d.n)
    V, alpha, beta = oracle_lib.Oracle(d).lanczos(x)
    assert alpha[-1] == 0.0
    G = V @ V.T
    assert np.abs(G - np.eye(10)).max() < 1e-12
    Vn, Tn, bn = onp.lanczos(onp.lap2d_uniform(50, 50, d.dx, d.dy), x, 10)
    assert rel_l2(V.T, Vn) < 1e-10 and abs(beta[0] - bn) < 1e-12 * bn


def test_batch_is_independent(oracle_lib):
    n, L = 16, 5.0
    d1 = Desc.make(n, n, L=L, equation=EQ.NLSE_CUBIC)
    d2 = Desc.make(n, n, L=L, equation=EQ.NLSE_CUBIC, batch=2)
    ua, ub = ics.gaussian_bumps_2d(n, L, seed=1), ics.gaussian_bumps_2d(n, L, seed=2)
    o2 = oracle_lib.Oracle(d2)
    o2.set(FIELD.U, np.stack([ua, ub])); o2.set(FIELD.M, np.ones(d1.n)); o2.step(1e-3, 3)
    o1 = oracle_lib.Oracle(d1)
    o1.set(FIELD.U, ub[None]); o1.set(FIELD.M, np.ones(d1.n)); o1.step(1e-3, 3)
    assert np.array_equal(o2.get()[1], o1.get()[0])


@pytest.mark.parametrize("eq,kind,host2d", [(EQ.NLSE_CUBIC, "cubic", True), (EQ.NLSE_CUBIC, "cubic", False),
                                            (EQ.NLSE_CQ, "cq", False), (EQ.NLSE_SAT, "sat", False)])
def test_sewi_steps_vs_numpy_twin(oracle_lib, eq, kind, host2d):
    """sEWI two-step integrator: SS2 start-up step, then sEWI steps (nlse_cubic_gautschi_driver.cpp:122-137)."""
    import nls_b200
    n, L, dt = 20, 5.0, 1e-3
    d = Desc.make(n, n, L=L, equation=eq, integrator=INT.SEWI, krylov_m=10, s1=1.0, s2=-0.1, kappa=0.3,
                  flags=nls_b200.FLAG_SEWI_HOST2D if host2d else 0)
    o = oracle_lib.Oracle(d)
    u0 = ics.gaussian_bumps_2d(n, L)
    mf = 1.0 + 0.1 * np.random.default_rng(4).standard_normal(d.n)
    o.set(FIELD.U, u0)
    o.set(FIELD.M, mf)
    o.step(dt, 5)
    A = onp.lap2d_uniform(n, n, d.dx, d.dy).astype(complex)
    u = u0.ravel().copy()
    uprev = u.copy()
    u = onp.neumann_bc(onp.nlse_step(A, u, mf, dt, 10, kind, 1.0, -0.1, 0.3).reshape(n, n)).ravel()
    for _ in range(4):
        un, uprev = onp.sewi_step(A, u, uprev, mf, dt, 10, kind, 1.0, -0.1, 0.3, host2d)
        u = onp.neumann_bc(un.reshape(n, n)).ravel()
    assert rel_l2(o.get()[0], u.reshape(n, n)) < 1e-12
    assert rel_l2(o.get(FIELD.U_PAST)[0], uprev.reshape(n, n)) < 1e-12
