"""Snapshot egress with on-device downsampling: nls_get_field_resampled.

The launcher downsamples every trajectory with `downsample_interpolation`
(finalized_scripts/downsampling.py:68-104; 3D :215-258; called from complex_launcher_2d.py:176-191)
AFTER the full-resolution snapshots crossed PCIe and the file system.  CPU: the oracle's restatement
against scipy's RegularGridInterpolator called exactly as the script calls it.  GPU: CUDA vs oracle.
"""
import numpy as np
import pytest
from scipy.interpolate import RegularGridInterpolator

import ics
from nls_b200 import Desc, EQ, INT, FIELD


def script_downsample_2d(u, target_shape, Lx, Ly):
    """downsampling.py:68-104 for one snapshot `u` of shape (n0, n1)."""
    n0, n1 = u.shape
    t0, t1 = target_shape
    x, y = np.linspace(-Lx, Lx, n0), np.linspace(-Ly, Ly, n1)
    X, Y = np.meshgrid(np.linspace(-Lx, Lx, t0), np.linspace(-Ly, Ly, t1), indexing='ij')
    f = RegularGridInterpolator((x, y), u, method='linear', bounds_error=False, fill_value=None)
    return f(np.vstack([X.ravel(), Y.ravel()]).T).reshape(t0, t1)


def script_downsample_3d(u, target_shape, L):
    """downsampling.py:215-258 for one snapshot."""
    grids = [np.linspace(-L, L, n) for n in u.shape]
    new = np.meshgrid(*[np.linspace(-L, L, t) for t in target_shape], indexing='ij')
    f = RegularGridInterpolator(tuple(grids), u, method='linear', bounds_error=False, fill_value=None)
    return f(np.vstack([g.ravel() for g in new]).T).reshape(target_shape)


def close(a, b, tol=1e-13):
    return np.max(np.abs(a - b)) <= tol * max(1.0, np.max(np.abs(b)))


@pytest.mark.parametrize("n,t", [(64, 16), (50, 17), (41, 41), (33, 50)])
def test_oracle_resample_matches_script_2d(oracle_lib, n, t):
    L = 5.0
    d = Desc.make(n, n, L=L, equation=EQ.NLSE_CUBIC, batch=2)
    rng = np.random.default_rng(n + t)
    u = rng.standard_normal((2, n, n)) + 1j * rng.standard_normal((2, n, n))
    o = oracle_lib.Oracle(d)
    o.set(FIELD.U, u)
    got = o.get_resampled(t, t)
    assert got.shape == (2, t, t)
    for b in range(2):
        assert close(got[b], script_downsample_2d(u[b], (t, t), L, L))


def test_oracle_resample_rectangular_and_real(oracle_lib):
    nx, ny, L, dt = 48, 36, 4.0, 1e-2
    d = Desc.make(nx, ny, L=L, equation=EQ.SGE, integrator=INT.GAUTSCHI, dt=dt)
    rng = np.random.default_rng(2)
    u, v = rng.standard_normal((ny, nx)), rng.standard_normal((ny, nx))
    o = oracle_lib.Oracle(d)
    o.set(FIELD.U, u)
    o.set(FIELD.V, v)
    assert close(o.get_resampled(20, 12)[0], script_downsample_2d(u, (12, 20), L, L))
    vv = o.get(FIELD.V)[0]
    assert close(o.get_resampled(20, 12, field=FIELD.V)[0], script_downsample_2d(vv, (12, 20), L, L), 1e-11)


def test_oracle_resample_matches_script_3d(oracle_lib):
    n, t, L = 20, 8, 3.0
    d = Desc.make(n, n, n, L=L, equation=EQ.NLSE_CUBIC)
    u = ics.gaussian_bumps_3d(n, L).reshape(n, n, n)
    o = oracle_lib.Oracle(d)
    o.set(FIELD.U, u)
    assert close(o.get_resampled(t, t, t)[0], script_downsample_3d(u, (t, t, t), L))


def test_oracle_resample_bad_args(oracle_lib):
    o = oracle_lib.Oracle(Desc.make(16, 16, L=1.0, equation=EQ.NLSE_CUBIC))
    with pytest.raises(Exception):
        o.get_resampled(1, 8)
    with pytest.raises(Exception):
        o.get_resampled(8, 8, 4)  # tnz on a 2D handle


# ------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("shape,target,batch", [((256, 256, 1), (128, 128, 1), 3), ((100, 70, 1), (33, 21, 1), 2),
                                                 ((40, 40, 40), (16, 16, 16), 1), ((24, 30, 18), (9, 11, 7), 2)])
def test_gpu_resample_complex(oracle_lib, shape, target, batch):
    from nls_b200 import Engine
    d = Desc.make(*shape, L=4.0, equation=EQ.NLSE_CUBIC, batch=batch)
    rng = np.random.default_rng(8)
    u = rng.standard_normal((batch, d.n)) + 1j * rng.standard_normal((batch, d.n))
    e, o = Engine(d), oracle_lib.Oracle(d)
    e.set(FIELD.U, u)
    o.set(FIELD.U, u)
    assert close(e.get_resampled(*target), o.get_resampled(*target), 1e-14)


@pytest.mark.gpu
def test_gpu_resample_real_after_steps(oracle_lib):
    """u and velocity snapshots of a running sine-Gordon trajectory, downsampled on the device."""
    from nls_b200 import Engine
    n, L, dt = 96, 6.0, 5e-3
    d = Desc.make(n, n, L=L, equation=EQ.SGE, integrator=INT.GAUTSCHI, krylov_m=10, dt=dt)
    e, o = Engine(d), oracle_lib.Oracle(d)
    for s in (e, o):
        s.set(FIELD.U, ics.kink_2d(n, L))
        s.set(FIELD.V, np.zeros(d.n))
        s.set(FIELD.M, np.ones(d.n))
        s.step(dt, 25)
    assert close(e.get_resampled(32, 32), o.get_resampled(32, 32), 1e-11)
    assert close(e.get_resampled(32, 32, field=FIELD.V), o.get_resampled(32, 32, field=FIELD.V), 1e-8)
    # identity when target == source
    assert close(e.get_resampled(n, n)[0], e.get()[0], 1e-14)
    # the state itself is untouched by the egress path
    e.step(dt, 5); o.step(dt, 5)
    assert np.linalg.norm(e.get() - o.get()) / np.linalg.norm(o.get()) < 1e-11


@pytest.mark.gpu
def test_gpu_resample_bad_args():
    from nls_b200 import Engine
    e = Engine(Desc.make(16, 16, L=1.0, equation=EQ.NLSE_CUBIC))
    for args in [(1, 8, 1), (8, 8, 4)]:
        with pytest.raises(Exception):
            e.get_resampled(*args)
    with pytest.raises(Exception):
        e.get_resampled(8, 8, field=FIELD.V)


# ---- spectral downsampling: downsample_fft (downsampling.py:10-34) as a direct cropped DFT on the device ------------
def test_fft_restatement_matches_the_script():
    """CPU, where /root/reference exists: the numpy restatement against the script's own torch.fft path"""
    import os
    import sys
    import types
    ref = "/root/reference/nlsolvers/finalized_scripts"
    if not os.path.exists(ref):
        pytest.skip("reference tree not present (GPU box)")
    import oracle_np
    mpl = types.ModuleType("matplotlib"); mpl.pyplot = types.ModuleType("matplotlib.pyplot")
    sys.modules.setdefault("matplotlib", mpl); sys.modules.setdefault("matplotlib.pyplot", mpl.pyplot)
    sys.path.insert(0, ref)
    import downsampling
    rng = np.random.default_rng(0)
    for shape, tgt in (((3, 64, 64), (16, 16)), ((2, 50, 37), (17, 12)), ((1, 33, 40), (33, 8))):
        for cplx in (True, False):
            u = rng.standard_normal(shape) + (1j * rng.standard_normal(shape) if cplx else 0)
            assert close(oracle_np.downsample_fft(u, tgt), downsampling.downsample_fft(u, tgt), 1e-13)


@pytest.mark.gpu
@pytest.mark.parametrize("ny,nx,ty,tx", [(64, 64, 16, 16), (50, 37, 17, 12), (33, 40, 33, 8), (256, 256, 128, 128)])
@pytest.mark.parametrize("kind", ["nlse", "sge"])
def test_gpu_fft_resample(ny, nx, ty, tx, kind):
    import nls_b200
    import oracle_np
    B, dt = 2, 1e-3
    rng = np.random.default_rng(5)
    if kind == "nlse":
        d = Desc.make(nx, ny, dx=0.1, equation=EQ.NLSE_CUBIC, batch=B)
        u0 = rng.standard_normal((B, ny, nx)) + 1j * rng.standard_normal((B, ny, nx))
    else:
        d = Desc.make(nx, ny, dx=0.1, equation=EQ.SGE, integrator=INT.SV, dt=dt, batch=B)
        u0 = rng.standard_normal((B, ny, nx))
    e = nls_b200.Engine(d)
    e.set(FIELD.U, u0)
    if kind == "sge":
        v0 = rng.standard_normal((B, ny, nx))
        e.set(FIELD.V, v0)
    got = e.get_fft_resampled(tx, ty)
    assert got.shape == (B, ty, tx)
    assert close(got, oracle_np.downsample_fft(u0, (ty, tx)), 1e-12)
    if kind == "sge":
        assert close(e.get_fft_resampled(tx, ty, field=FIELD.V), oracle_np.downsample_fft(e.get(FIELD.V), (ty, tx)), 1e-12)
    # a second target shape on the same handle rebuilds the weights
    assert close(e.get_fft_resampled(max(tx // 2, 1), ty), oracle_np.downsample_fft(u0, (ty, max(tx // 2, 1))), 1e-12)
