"""The reference's driver command lines on the B200 engine.

CPU part: argv contract, usage/exit codes, shape checks, loud failure without a GPU.
GPU part (-m gpu): whole trajectories written by the CLIs vs the oracle driven through the same
loop semantics as the reference drivers (nlse_cubic_driver.cpp:96-115, sg_single_driver.cpp:110-124).
"""
import os
import subprocess

import numpy as np
import pytest

from conftest import rel_l2
import ics
import nls_b200
from nls_b200 import Desc, EQ, INT, LAP, FIELD

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "nonlinear-solvers_b200", "host")
BIN = os.path.join(HOST, "bin")


@pytest.fixture(scope="module")
def drivers():
    nls_b200.build_library()
    subprocess.run(["make", "-C", HOST, "-s"], check=True)
    return BIN


def run(drivers, name, *args, check=False):
    return subprocess.run([os.path.join(drivers, name)] + [str(a) for a in args], capture_output=True, text=True,
                          check=check)


def test_usage_on_wrong_argc(drivers):
    for name in ["nlse_cubic", "nlse_cubic_quintic", "nlse_saturating", "sg_single", "kg_sv", "nlse_3d_dev",
                 "kg_gautschi_3d_dev", "nlse_2d_dev"]:
        r = run(drivers, name, 16, 16)
        assert r.returncode == 1 and "Usage:" in r.stderr, name
    r = run(drivers, "nls_driver", "no_such_driver")
    assert r.returncode == 1 and "Known drivers" in r.stderr


def test_shape_mismatch_returns_1(drivers, tmp_path):
    np.save(tmp_path / "u0.npy", np.zeros((8, 9), dtype=np.complex128))
    r = run(drivers, "nlse_cubic", 16, 16, 5.0, 5.0, tmp_path / "u0.npy", tmp_path / "o.npy", 0.01, 10, 5)
    assert r.returncode == 1 and "dimensions mismatch" in r.stderr


def test_wrong_dtype_aborts(drivers, tmp_path):
    np.save(tmp_path / "u0.npy", np.zeros((16, 16), dtype=np.float32))
    r = run(drivers, "nlse_cubic", 16, 16, 5.0, 5.0, tmp_path / "u0.npy", tmp_path / "o.npy", 0.01, 10, 5)
    assert r.returncode != 0 and "dtype" in r.stderr


def test_no_gpu_fails_loudly(drivers, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    np.save(tmp_path / "u0.npy", ics.gaussian_bumps_2d(16, 5.0))
    r = run(drivers, "nlse_cubic", 16, 16, 5.0, 5.0, tmp_path / "u0.npy", tmp_path / "o.npy", 0.01, 10, 5)
    assert r.returncode != 0 and not (tmp_path / "o.npy").exists()
    assert "nls_create" in r.stderr


# ------------------------------------------------------------------------------------- GPU
def oracle_trajectory(oracle_lib, d, u0, v0, m, c, T, nt, ns):
    """The reference driver loop on the oracle."""
    dt, freq = T / nt, nt // ns
    o = oracle_lib.Oracle(d)
    if c is not None:
        o.set(FIELD.C, c)
    o.set(FIELD.M, m)
    o.set(FIELD.U, u0)
    if v0 is not None:
        o.set(FIELD.V, v0)
    tu = np.zeros((ns,) + u0.shape, dtype=u0.dtype)
    tv = np.zeros_like(tu)
    tu[0] = u0
    if v0 is not None:
        tv[0] = v0
    for i in range(1, nt):
        o.step(dt, 1)
        if i % freq == 0 and i // freq < ns:
            tu[i // freq] = o.get()[0]
            if v0 is not None:
                tv[i // freq] = o.get(FIELD.V)[0]
    return tu, tv


@pytest.mark.gpu
@pytest.mark.parametrize("name,eq,extra,mdef", [
    ("nlse_cubic", EQ.NLSE_CUBIC, [], None), ("nlse_cubic", EQ.NLSE_CUBIC, [], 1.0),
    ("nlse_cubic_quintic", EQ.NLSE_CQ, [1.0, -0.1], 1.0), ("nlse_saturating", EQ.NLSE_SAT, [0.3], 1.0)])
def test_nlse_cli_trajectory(drivers, oracle_lib, tmp_path, name, eq, extra, mdef):
    n, L, T, nt, ns = 48, 6.0, 0.05, 50, 10
    u0 = ics.gaussian_bumps_2d(n, L)
    np.save(tmp_path / "u0.npy", u0)
    args = [n, n, L, L] + extra + [tmp_path / "u0.npy", tmp_path / "traj.npy", T, nt, ns]
    m = np.zeros(n * n)  # driver default m == 0 (nlse_cubic_driver.cpp:63)
    if mdef is not None:
        m = np.full(n * n, mdef) * (1 + 0.1 * np.cos(np.arange(n * n)))
        np.save(tmp_path / "m.npy", m.reshape(n, n))
        args.append(tmp_path / "m.npy")
    run(drivers, name, *args, check=True)
    got = np.load(tmp_path / "traj.npy")
    assert got.shape == (ns, n, n) and got.dtype == np.complex128
    s1, s2 = (extra + [0, 0])[:2]
    d = Desc.make(n, n, L=L, equation=eq, krylov_m=10, s1=s1, s2=s2, kappa=s1)
    ref, _ = oracle_trajectory(oracle_lib, d, u0, None, m, None, T, nt, ns)
    assert np.array_equal(got[0], u0)
    assert rel_l2(got, ref) < 1e-11


@pytest.mark.gpu
@pytest.mark.parametrize("name,eq,integ,sign", [
    ("sg_single", EQ.SGE, INT.GAUTSCHI, 1.0), ("sg_single_sv", EQ.SGE, INT.SV, 1.0), ("kg", EQ.KGE, INT.GAUTSCHI, -1.0),
    ("kg_sv", EQ.KGE, INT.SV, 1.0), ("phi4", EQ.PHI4, INT.GAUTSCHI, -1.0), ("sg_double", EQ.SGE_DOUBLE, INT.GAUTSCHI, -1.0),
    ("sg_hyperbolic", EQ.SGE_SINH, INT.GAUTSCHI, -1.0)])
def test_real_cli_trajectory(drivers, oracle_lib, tmp_path, name, eq, integ, sign):
    n, L, T, nt, ns = 48, 6.0, 0.2, 40, 8
    u0 = 0.3 * ics.kink_2d(n, L)
    v0 = 0.05 * np.cos(u0)
    np.save(tmp_path / "u0.npy", u0)
    np.save(tmp_path / "v0.npy", v0)
    m = 1.0 + 0.2 * np.sin(np.arange(n * n) / 7.0)
    np.save(tmp_path / "m.npy", m.reshape(n, n))
    run(drivers, name, n, n, L, L, tmp_path / "u0.npy", tmp_path / "v0.npy", tmp_path / "tu.npy", tmp_path / "tv.npy",
        T, nt, ns, tmp_path / "m.npy", check=True)
    d = Desc.make(n, n, L=L, equation=eq, integrator=integ, krylov_m=10, dt=T / nt, op_sign=sign)
    ru, rv = oracle_trajectory(oracle_lib, d, u0, v0, m, None, T, nt, ns)
    gu, gv = np.load(tmp_path / "tu.npy"), np.load(tmp_path / "tv.npy")
    assert gu.shape == (ns, n, n)
    assert rel_l2(gu, ru) < 1e-11
    assert rel_l2(gv, rv) < 1e-8  # (u - u_past)/dt amplifies rounding by 1/dt


@pytest.mark.gpu
def test_3d_device_driver_cli(drivers, oracle_lib, tmp_path):
    n, L, T, nt, ns = 20, 3.0, 0.02, 20, 5
    u0 = ics.gaussian_bumps_3d(n, L)
    rng = np.random.default_rng(11)
    m = np.ones(n ** 3)
    c = rng.uniform(0.8, 1.2, n ** 3)
    for nm, a in [("u0", u0), ("m", m.reshape(n, n, n)), ("c", c.reshape(n, n, n))]:
        np.save(tmp_path / f"{nm}.npy", a)
    run(drivers, "nlse_3d_dev", n, n, n, L, L, L, tmp_path / "u0.npy", tmp_path / "traj.npy", T, nt, ns,
        tmp_path / "m.npy", tmp_path / "c.npy", check=True)
    d = Desc.make(n, n, n, L=L, equation=EQ.NLSE_CUBIC, laplacian=LAP.ANISO, krylov_m=25)
    ref, _ = oracle_trajectory(oracle_lib, d, u0, None, m, c, T, nt, ns)
    got = np.load(tmp_path / "traj.npy")
    assert got.shape == (ns, n, n, n)
    assert rel_l2(got, ref) < 1e-11


# ---------------------------------------------------------------- ensemble files
def test_ensemble_velocity_count_mismatch_returns_1(drivers, tmp_path):
    """u0 and v0 must hold the same number of trajectories; checked before any CUDA call."""
    np.save(tmp_path / "u0.npy", np.zeros((3, 16, 16)))
    np.save(tmp_path / "v0.npy", np.zeros((2, 16, 16)))
    r = run(drivers, "sg_single", 16, 16, 5.0, 5.0, tmp_path / "u0.npy", tmp_path / "v0.npy", tmp_path / "tu.npy",
            tmp_path / "tv.npy", 0.01, 10, 5)
    assert r.returncode == 1 and "different number of trajectories" in r.stderr


def test_ensemble_bad_inner_shape_returns_1(drivers, tmp_path):
    np.save(tmp_path / "u0.npy", np.zeros((3, 16, 15), dtype=np.complex128))
    r = run(drivers, "nlse_ensemble_driver", 16, 16, 5.0, 5.0, tmp_path / "u0.npy", tmp_path / "o.npy", 0.01, 10, 5)
    assert r.returncode == 1 and "dimensions mismatch" in r.stderr


def test_ensemble_too_many_devices_returns_1(drivers, tmp_path):
    np.save(tmp_path / "u0.npy", np.zeros((3, 16, 16), dtype=np.complex128))
    r = run(drivers, "nlse_ensemble_driver", 16, 16, 5.0, 5.0, tmp_path / "u0.npy", tmp_path / "o.npy", 0.01, 10, 5,
            "--devices", 64)
    assert r.returncode == 1 and "CUDA device(s) are visible" in r.stderr and not (tmp_path / "o.npy").exists()


def random_bumps(B, n, L, seed=7):
    """B distinct smooth complex ICs: the sampler's random-IC ensemble at test size."""
    rng = np.random.default_rng(seed)
    x = np.linspace(-L, L, n)
    X, Y = np.meshgrid(x, x)
    out = np.empty((B, n, n), dtype=np.complex128)
    for b in range(B):
        cx, cy, kx, ky = rng.uniform(-2, 2, 4)
        out[b] = np.exp(-((X - cx) ** 2 + (Y - cy) ** 2)) * np.exp(1j * (kx * X + ky * Y))
    return out


@pytest.mark.gpu
def test_nlse_ensemble_cli(drivers, oracle_lib, tmp_path):
    """One (B,ny,nx) file -> (B,ns,ny,nx): every trajectory equals the oracle's and, bit for bit,
    what the single-trajectory driver writes for the same initial condition."""
    B, n, L, T, nt, ns = 5, 40, 5.0, 0.04, 40, 8
    u0 = random_bumps(B, n, L)
    m = 1.0 + 0.1 * np.cos(np.arange(n * n) / 3.0)
    np.save(tmp_path / "u0.npy", u0)
    np.save(tmp_path / "m.npy", m.reshape(n, n))
    run(drivers, "nlse_ensemble_driver", n, n, L, L, tmp_path / "u0.npy", tmp_path / "traj.npy", T, nt, ns,
        tmp_path / "m.npy", check=True)
    got = np.load(tmp_path / "traj.npy")
    assert got.shape == (B, ns, n, n) and got.dtype == np.complex128
    d = Desc.make(n, n, L=L, equation=EQ.NLSE_CUBIC, krylov_m=10)
    for b in range(B):
        ref, _ = oracle_trajectory(oracle_lib, d, u0[b], None, m, None, T, nt, ns)
        assert np.array_equal(got[b, 0], u0[b])
        assert rel_l2(got[b], ref) < 1e-11, b
    for b in (0, B - 1):
        np.save(tmp_path / "one.npy", u0[b])
        run(drivers, "nlse_cubic", n, n, L, L, tmp_path / "one.npy", tmp_path / "one_traj.npy", T, nt, ns,
            tmp_path / "m.npy", check=True)
        assert rel_l2(got[b], np.load(tmp_path / "one_traj.npy")) < 1e-13
    # default m == 1 for the ensemble alias (nlse_cubic keeps the reference's m == 0)
    run(drivers, "nlse_ensemble_driver", n, n, L, L, tmp_path / "u0.npy", tmp_path / "t1.npy", T, nt, ns, check=True)
    ref, _ = oracle_trajectory(oracle_lib, d, u0[1], None, np.ones(n * n), None, T, nt, ns)
    assert rel_l2(np.load(tmp_path / "t1.npy")[1], ref) < 1e-11


@pytest.mark.gpu
def test_real_ensemble_cli(drivers, oracle_lib, tmp_path):
    """Ensemble files through a real-valued driver (sine-Gordon Gautschi): u and velocity trajectories."""
    B, n, L, T, nt, ns = 3, 40, 6.0, 0.2, 40, 8
    base = 0.3 * ics.kink_2d(n, L)
    u0 = np.stack([base * (1 + 0.2 * b) for b in range(B)])
    v0 = 0.05 * np.cos(u0)
    np.save(tmp_path / "u0.npy", u0)
    np.save(tmp_path / "v0.npy", v0)
    m = 1.0 + 0.2 * np.sin(np.arange(n * n) / 7.0)
    np.save(tmp_path / "m.npy", m.reshape(n, n))
    run(drivers, "sg_single", n, n, L, L, tmp_path / "u0.npy", tmp_path / "v0.npy", tmp_path / "tu.npy",
        tmp_path / "tv.npy", T, nt, ns, tmp_path / "m.npy", check=True)
    gu, gv = np.load(tmp_path / "tu.npy"), np.load(tmp_path / "tv.npy")
    assert gu.shape == (B, ns, n, n) and gv.shape == (B, ns, n, n)
    d = Desc.make(n, n, L=L, equation=EQ.SGE, integrator=INT.GAUTSCHI, krylov_m=10, dt=T / nt)
    for b in range(B):
        ru, rv = oracle_trajectory(oracle_lib, d, u0[b], v0[b], m, None, T, nt, ns)
        assert rel_l2(gu[b], ru) < 1e-11
        assert rel_l2(gv[b], rv) < 1e-8


@pytest.mark.gpu
def test_ensemble_cli_device_split(drivers, tmp_path):
    """--devices G block-partitions the trajectories; the result does not depend on G."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    B, n, L, T, nt, ns = 5, 40, 5.0, 0.04, 40, 8
    np.save(tmp_path / "u0.npy", random_bumps(B, n, L))
    outs = []
    for G in (1, 2):
        run(drivers, "nlse_ensemble_driver", n, n, L, L, tmp_path / "u0.npy", tmp_path / f"t{G}.npy", T, nt, ns,
            "--devices", G, check=True)
        outs.append(np.load(tmp_path / f"t{G}.npy"))
    assert rel_l2(outs[1], outs[0]) < 1e-13


@pytest.mark.gpu
def test_cli_diagnostics_file(drivers, tmp_path):
    """--diag: mass and energy of every stored snapshot equal the reference scripts' formulas applied
    to the trajectory files (compare_utils_complex_2d.py:134-151, energy_sge.py:6-14)."""
    from test_energy import nlse_hamiltonian_script, sge_energy_script
    B, n, L, T, nt, ns = 3, 40, 5.0, 0.04, 40, 8
    dx = 2 * L / (n - 1)
    np.save(tmp_path / "u0.npy", random_bumps(B, n, L))
    run(drivers, "nlse_ensemble_driver", n, n, L, L, tmp_path / "u0.npy", tmp_path / "t.npy", T, nt, ns,
        "--diag", tmp_path / "d.npy", check=True)
    tr, dg = np.load(tmp_path / "t.npy"), np.load(tmp_path / "d.npy")
    assert dg.shape == (B, ns, 2)
    for b in range(B):
        for s in range(ns):
            assert dg[b, s, 0] == pytest.approx(np.sum(np.abs(tr[b, s]) ** 2) * dx * dx, rel=1e-12)
            assert dg[b, s, 1] == pytest.approx(nlse_hamiltonian_script(tr[b, s], dx, dx), rel=1e-11)
    assert np.max(np.abs(dg[:, :, 0] / dg[:, :1, 0] - 1)) < 1e-6  # SS2 conserves mass up to the BC copy

    u0 = 0.3 * ics.kink_2d(n, L)
    np.save(tmp_path / "ru.npy", u0)
    np.save(tmp_path / "rv.npy", 0.05 * np.cos(u0))
    run(drivers, "sg_single", n, n, L, L, tmp_path / "ru.npy", tmp_path / "rv.npy", tmp_path / "tu.npy",
        tmp_path / "tv.npy", 0.2, 40, 8, "--diag", tmp_path / "rd.npy", check=True)
    tu, tv, rd = np.load(tmp_path / "tu.npy"), np.load(tmp_path / "tv.npy"), np.load(tmp_path / "rd.npy")
    assert rd.shape == (8, 2)
    for s in range(8):
        assert rd[s, 1] == pytest.approx(sge_energy_script(tu[s], tv[s], dx, dx), rel=1e-9)


def test_downsample_flag_validation(drivers, tmp_path):
    np.save(tmp_path / "u0.npy", np.zeros((16, 16), dtype=np.complex128))
    r = run(drivers, "nlse_cubic", 16, 16, 5.0, 5.0, tmp_path / "u0.npy", tmp_path / "o.npy", 0.01, 10, 5,
            "--downsample", 1)
    assert r.returncode == 1 and "--downsample" in r.stderr


@pytest.mark.gpu
def test_cli_downsampled_trajectory(drivers, tmp_path):
    """--downsample N stores what the launcher's downsample_interpolation makes of the full-resolution
    trajectory (complex_launcher_2d.py:176-191 -> downsampling.py:68-104), for ensemble and real drivers."""
    from test_downsample import script_downsample_2d
    B, n, N, L, T, nt, ns = 3, 48, 16, 5.0, 0.04, 40, 8
    np.save(tmp_path / "u0.npy", random_bumps(B, n, L))
    common = [n, n, L, L, tmp_path / "u0.npy"]
    run(drivers, "nlse_ensemble_driver", *common, tmp_path / "full.npy", T, nt, ns, check=True)
    run(drivers, "nlse_ensemble_driver", *common, tmp_path / "small.npy", T, nt, ns, "--downsample", N, check=True)
    full, small = np.load(tmp_path / "full.npy"), np.load(tmp_path / "small.npy")
    assert small.shape == (B, ns, N, N) and small.dtype == np.complex128
    for b in range(B):
        for s in range(ns):
            ref = script_downsample_2d(full[b, s], (N, N), L, L)
            assert np.max(np.abs(small[b, s] - ref)) < 1e-13 * max(1.0, np.max(np.abs(ref)))

    u0 = 0.3 * ics.kink_2d(n, L)
    np.save(tmp_path / "ru.npy", u0)
    np.save(tmp_path / "rv.npy", 0.05 * np.cos(u0))
    rc = [n, n, L, L, tmp_path / "ru.npy", tmp_path / "rv.npy"]
    run(drivers, "sg_single", *rc, tmp_path / "fu.npy", tmp_path / "fv.npy", 0.2, 40, 8, check=True)
    run(drivers, "sg_single", *rc, tmp_path / "su.npy", tmp_path / "sv.npy", 0.2, 40, 8, "--downsample", N, check=True)
    fu, fv = np.load(tmp_path / "fu.npy"), np.load(tmp_path / "fv.npy")
    su, sv = np.load(tmp_path / "su.npy"), np.load(tmp_path / "sv.npy")
    assert su.shape == (8, N, N) and sv.shape == (8, N, N)
    for s in range(8):
        assert np.max(np.abs(su[s] - script_downsample_2d(fu[s], (N, N), L, L))) < 1e-12
        assert np.max(np.abs(sv[s] - script_downsample_2d(fv[s], (N, N), L, L))) < 1e-9
    # --downsample-fft N: the launcher's downsample_fft (downsampling.py:10-34) of the full-resolution trajectory
    import oracle_np
    run(drivers, "nlse_ensemble_driver", *common, tmp_path / "fft.npy", T, nt, ns, "--downsample-fft", N, check=True)
    sf = np.load(tmp_path / "fft.npy")
    assert sf.shape == (B, ns, N, N)
    for b in range(B):
        ref = oracle_np.downsample_fft(full[b], (N, N))
        assert np.max(np.abs(sf[b] - ref)) < 1e-12 * max(1.0, np.max(np.abs(ref)))
    run(drivers, "sg_single", *rc, tmp_path / "gu.npy", tmp_path / "gv.npy", 0.2, 40, 8, "--downsample-fft", N, check=True)
    assert np.max(np.abs(np.load(tmp_path / "gu.npy") - oracle_np.downsample_fft(fu, (N, N)))) < 1e-12
