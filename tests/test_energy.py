"""NLS_RED_ENERGY: the energies of the reference's drift scripts (SURVEY.md 8c "mass / energy
definitions", 8f N4), fused into one reduction pass.

CPU: the oracle's restatement against the scripts' own numpy expressions, restated here line by
line (energy_sge.py:6-14, energy_nlse.py is NOT the launcher's definition;
finalized_scripts/compare_utils_complex_2d.py:134-151 is).
GPU: the CUDA reduction against the oracle, and the energy series of whole trajectories.
"""
import numpy as np
import pytest

import ics
from nls_b200 import Desc, EQ, INT, FIELD, RED


def sge_energy_script(u, v, dx, dy):
    """energy_sge.py:6-14, verbatim arithmetic."""
    u_x = np.gradient(u, dx, axis=0)
    u_y = np.gradient(u, dy, axis=1)
    t = np.sum(.5 * (u_x ** 2 + u_y ** 2 + v ** 2 + 2 * (1 - np.cos(u))))
    return .5 * dx * dy * t


def nlse_hamiltonian_script(u, dx, dy):
    """compare_utils_complex_2d.py:134-151: gradient part + potential part."""
    ux = (u[1:-1, 2:] - u[1:-1, :-2]) / (2. * dx)
    uy = (u[2:, 1:-1] - u[:-2, 1:-1]) / (2. * dy)
    grad = np.sum(np.abs(ux) ** 2 + np.abs(uy) ** 2) * dx * dy
    pot = np.sum(-0.5 * np.abs(u) ** 4) * dx * dy
    return grad + pot


def test_oracle_energy_matches_scripts_nlse(oracle_lib):
    nx, ny, L = 40, 40, 6.0
    d = Desc.make(nx, ny, L=L, equation=EQ.NLSE_CUBIC, batch=2)
    u = np.stack([ics.gaussian_bumps_2d(nx, L, seed=s) for s in (1, 2)])
    o = oracle_lib.Oracle(d)
    o.set(FIELD.U, u)
    got = o.reduce(RED.ENERGY)
    for b in range(2):
        assert got[b] == pytest.approx(nlse_hamiltonian_script(u[b].reshape(ny, nx), d.dx, d.dy), rel=1e-13)


def test_oracle_energy_matches_scripts_sge(oracle_lib):
    n, L, dt = 36, 6.0, 5e-3
    d = Desc.make(n, n, L=L, equation=EQ.SGE, integrator=INT.GAUTSCHI, dt=dt)
    d.dy = 1.25 * d.dx  # the script's axis/spacing pairing only shows when dx != dy
    u = 0.7 * ics.kink_2d(n, L)
    v = 0.1 * np.sin(3 * u) + 0.05
    o = oracle_lib.Oracle(d)
    o.set(FIELD.U, u)
    o.set(FIELD.V, v)
    vv = o.get(FIELD.V)[0].reshape(n, n)  # (u - u_past)/dt, what the driver stores (sg_single_driver.cpp:113)
    assert o.reduce(RED.ENERGY)[0] == pytest.approx(sge_energy_script(u.reshape(n, n), vv, d.dx, d.dy), rel=1e-13)


def test_oracle_energy_3d_plane_wave(oracle_lib):
    """3D extension: a field that varies along z only has |u_z|^2 as its whole gradient energy."""
    n, L = 12, 2.0
    d = Desc.make(n, n, n, L=L, equation=EQ.NLSE_CUBIC)
    z = np.linspace(-L, L, n)
    u = np.broadcast_to(np.exp(1j * 0.7 * z)[:, None, None], (n, n, n)).copy()
    o = oracle_lib.Oracle(d)
    o.set(FIELD.U, u)
    grad = (np.sin(0.7 * d.dz) / d.dz) ** 2 * (n - 2) ** 3 * d.dx * d.dy * d.dz
    pot = -0.5 * n ** 3 * d.dx * d.dy * d.dz
    assert o.reduce(RED.ENERGY)[0] == pytest.approx(grad + pot, rel=1e-12)


# ------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(48, 48, 1), (33, 47, 1), (20, 20, 20), (9, 14, 11)])
def test_gpu_energy_complex(oracle_lib, shape):
    from nls_b200 import Engine
    nx, ny, nz = shape
    rng = np.random.default_rng(5)
    d = Desc.make(nx, ny, nz, L=4.0, equation=EQ.NLSE_CUBIC, batch=3)
    d.dy, d.dz = 1.25 * d.dx, 0.8 * d.dx
    u = rng.standard_normal((3, d.n)) + 1j * rng.standard_normal((3, d.n))
    e, o = Engine(d), oracle_lib.Oracle(d)
    e.set(FIELD.U, u)
    o.set(FIELD.U, u)
    assert np.allclose(e.reduce(RED.ENERGY), o.reduce(RED.ENERGY), rtol=1e-12, atol=0)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(48, 48, 1), (33, 47, 1), (16, 16, 16), (9, 14, 11)])
def test_gpu_energy_real(oracle_lib, shape):
    from nls_b200 import Engine
    nx, ny, nz = shape
    rng = np.random.default_rng(6)
    d = Desc.make(nx, ny, nz, L=4.0, equation=EQ.SGE, integrator=INT.GAUTSCHI, dt=1e-2, batch=2)
    d.dy, d.dz = 1.25 * d.dx, 0.8 * d.dx
    u, v = rng.standard_normal((2, d.n)), rng.standard_normal((2, d.n))
    e, o = Engine(d), oracle_lib.Oracle(d)
    for s in (e, o):
        s.set(FIELD.U, u)
        s.set(FIELD.V, v)
    assert np.allclose(e.reduce(RED.ENERGY), o.reduce(RED.ENERGY), rtol=1e-12, atol=0)


@pytest.mark.gpu
def test_energy_series_matches_oracle(oracle_lib):
    """North star: energy drift of the engine's trajectory matches the reference path's to 1e-8.
    sine-Gordon kink, Gautschi, 10 snapshots of 20 steps; and NLSE SS2."""
    from nls_b200 import Engine
    n, L, dt = 64, 6.0, 5e-3
    d = Desc.make(n, n, L=L, equation=EQ.SGE, integrator=INT.GAUTSCHI, krylov_m=10, dt=dt)
    u0 = ics.kink_2d(n, L)
    e, o = Engine(d), oracle_lib.Oracle(d)
    for s in (e, o):
        s.set(FIELD.U, u0)
        s.set(FIELD.V, np.zeros(d.n))
        s.set(FIELD.M, np.ones(d.n))
    series = []
    for _ in range(10):
        e.step(dt, 20)
        o.step(dt, 20)
        series.append((e.reduce(RED.ENERGY)[0], o.reduce(RED.ENERGY)[0]))
    s = np.array(series)
    assert np.max(np.abs(s[:, 0] - s[:, 1]) / np.abs(s[:, 1])) < 1e-8
    # same quantity from the host copy through the script formula
    uu, vv = e.get()[0].reshape(n, n), e.get(FIELD.V)[0].reshape(n, n)
    assert s[-1, 0] == pytest.approx(sge_energy_script(uu, vv, d.dx, d.dy), rel=1e-12)

    dc = Desc.make(n, n, L=L, equation=EQ.NLSE_CUBIC, krylov_m=10)
    ec, oc = Engine(dc), oracle_lib.Oracle(dc)
    for s_ in (ec, oc):
        s_.set(FIELD.U, ics.gaussian_bumps_2d(n, L))
        s_.set(FIELD.M, np.ones(dc.n))
    for _ in range(5):
        ec.step(1e-3, 20)
        oc.step(1e-3, 20)
        a, b = ec.reduce(RED.ENERGY)[0], oc.reduce(RED.ENERGY)[0]
        assert abs(a - b) / abs(b) < 1e-8
    assert a == pytest.approx(nlse_hamiltonian_script(ec.get()[0].reshape(n, n), dc.dx, dc.dy), rel=1e-12)


@pytest.mark.gpu
def test_energy_needs_dt_for_real_state():
    from nls_b200 import Engine
    d = Desc.make(16, 16, L=4.0, equation=EQ.SGE, integrator=INT.SV)  # dt = 0 in the descriptor
    e = Engine(d)
    e.set(FIELD.U, np.zeros(d.n))
    with pytest.raises(Exception):
        e.reduce(RED.ENERGY)
