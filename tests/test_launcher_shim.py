"""tools/launcher_shim.py: the reference launchers' one-process-per-run loop handed to the engine
as ensemble batches.  CPU: grouping / ordering / the patched run() against a recording fake driver.  GPU: the
batched call writes bit-identical trajectories to the per-run calls."""
import os
import stat
import sys
import types

import numpy as np
import pytest

import ics

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import launcher_shim as shim  # noqa: E402

BIN = os.path.join(ROOT, "nonlinear-solvers_b200", "host", "bin")

FAKE = '''#!{py}
import sys, numpy as np
a = sys.argv[1:]
nx, ny = int(a[0]), int(a[1])
u0 = np.load(a[4]); ns = int(a[8])
m = np.load(a[9]) if len(a) > 9 and a[9].endswith(".npy") else np.zeros((ny, nx))
c = np.load(a[10]) if len(a) > 10 and a[10].endswith(".npy") else np.zeros((ny, nx))
ens = u0.ndim == 3
U = u0 if ens else u0[None]
traj = np.stack([np.stack([u * (s + 1) + m.sum() + 2 * c.sum() for s in range(ns)]) for u in U])
np.save(a[5], traj if ens else traj[0])
open(a[5] + ".calls", "a").write("x")
'''


@pytest.fixture
def fake_exe(tmp_path):
    p = tmp_path / "fake_driver"
    p.write_text(FAKE.format(py=sys.executable))
    p.chmod(p.stat().st_mode | stat.S_IEXEC)
    return str(p)


def test_grouping_by_shared_fields():
    m1, m2 = np.ones((4, 4)), 2 * np.ones((4, 4))
    assert shim.group_runs([m1, m2, m1.copy(), None], [None, None, None, None]) == [[0, 2], [1], [3]]


def test_run_batch_order_and_process_count(fake_exe, tmp_path):
    rng = np.random.default_rng(0)
    u0s = [rng.standard_normal((6, 5)) + 1j * rng.standard_normal((6, 5)) for _ in range(5)]
    m1, m2 = np.ones((6, 5)), 3 * np.ones((6, 5))
    ms = [m1, m2, m1, m1, m2]
    trajs, _ = shim.run_batch(fake_exe, 5, 6, 1.0, 1.0, u0s, ms, ms, 1.0, 10, 3, workdir=str(tmp_path))
    for u0, m, t in zip(u0s, ms, trajs):
        assert t.shape == (3, 6, 5)
        assert np.allclose(t[2], u0 * 3 + m.sum() + 2 * m.sum())


def test_patched_launcher_keeps_rng_order_and_c_is_m(fake_exe, tmp_path):
    class Launcher:  # the attribute / method names of complex_launcher_2d.NLSELauncher
        def __init__(self):
            self.args = types.SimpleNamespace(num_runs=4, exe=fake_exe, nx=5, ny=6, Lx=1.0, Ly=1.0, T=1.0, nt=10, snapshots=2)
            self.saved = []

        def generate_initial_condition(self, i):
            return np.random.standard_normal((6, 5)) + 0j, {"i": i}

        def generate_anisotropy(self, i):
            return np.full((6, 5), 7.0)

        def generate_spatial_amplification(self, i, c):
            return np.full((6, 5), float(i % 2))  # two distinct m fields -> two groups

        def downsample_trajectory(self, t):
            return t[:, ::2]

        def save_to_hdf5(self, i, u0, m, c, traj, params, wall):
            self.saved.append((i, traj, u0, m))

        def cleanup(self, i):
            pass

    shim.patch_launcher(Launcher)
    np.random.seed(5)
    la = Launcher()
    la.run()
    np.random.seed(5)
    expect_u0 = [np.random.standard_normal((6, 5)) for _ in range(4)]
    assert [s[0] for s in la.saved] == [0, 1, 2, 3]
    for i, traj, u0, m in la.saved:
        assert np.array_equal(u0.real, expect_u0[i])
        assert np.allclose(traj[1], (u0 * 2 + m.sum() + 2 * m.sum())[::2])  # c file == m (launcher quirk), not 7


@pytest.mark.gpu
def test_batched_launch_equals_per_run_launches(tmp_path):
    import nls_b200
    import subprocess
    nls_b200.build_library()
    subprocess.run(["make", "-C", os.path.join(ROOT, "nonlinear-solvers_b200", "host"), "-s"], check=True)
    n, L, T, nt, ns, B = 48, 6.0, 0.03, 30, 6, 5
    u0s = [ics.gaussian_bumps_2d(n, L, seed=10000 + t) for t in range(B)]
    m = 1.0 + 0.1 * np.cos(np.arange(n * n)).reshape(n, n)
    exe = os.path.join(BIN, "nlse_2d_dev")
    batched, _ = shim.run_batch(exe, n, n, L, L, u0s, [m] * B, [m] * B, T, nt, ns, workdir=str(tmp_path))
    for t in (0, 3):
        single, _ = shim.run_batch(exe, n, n, L, L, [u0s[t]], [m], [m], T, nt, ns, workdir=str(tmp_path))
        assert batched[t].shape == (ns, n, n)
        err = np.linalg.norm(batched[t] - single[0]) / np.linalg.norm(single[0])
        assert err < 1e-12, err  # batch handle (multi-kernel path) vs single handle (persistent kernel): rounding only
