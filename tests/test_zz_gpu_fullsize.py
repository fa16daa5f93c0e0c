"""Full-size GPU parity (run with -m gpu on a B200): every BASELINE config at the grid size and launch
shape its bench number comes from, CUDA path (through the C ABI) vs the CPU oracle, plus the cross-checks
between launch shapes that SURVEY.md 8d / VERDICT r01 ask for.  The oracle runs are started together in a
thread pool by tests/fullsize.py (conftest kicks it off when the GPU suite starts); this file sorts last so
they overlap with the rest of the suite.  Bar: rel-L2 <= 1e-10 (north star), worst errors printed."""
import os

import numpy as np
import pytest

from conftest import rel_l2
import fullsize
import ics
import nls_b200
from nls_b200 import Desc, Engine, EQ, INT, LAP, FIELD, RED

pytestmark = pytest.mark.gpu
CASES = fullsize.cases()


def compare(name, got, ref, tol_u=1e-10, tol_v=1e-6, members=None):
    worst = 0.0
    for cp in sorted(ref):
        gu = got[cp][0] if members is None else got[cp][0][list(members)]
        eu = rel_l2(gu, ref[cp][0])
        msg = f"{name}: step {cp}: rel-L2(u) = {eu:.3e}"
        assert np.isfinite(gu).all()
        if len(ref[cp]) > 1:
            ev = rel_l2(got[cp][1], ref[cp][1])
            msg += f", rel-L2(v) = {ev:.3e}"
            assert ev <= tol_v, msg
        print(msg)
        assert eu <= tol_u, msg
        worst = max(worst, eu)
    return worst


@pytest.mark.parametrize("name", ["c2_gautschi_1024", "c2_gautschi_256", "c2_sv_1024", "c2_sv_256"])
def test_config2_full_size(name):
    """BASELINE configs[1]: sine-Gordon 1024^2 (first 200 Gautschi steps / 2000 Stoermer-Verlet steps: the
    148-CTA persistent kernel with 7086-site chunks, the k_sv launch shape of the bench) and 256^2 for the
    whole 4999-step driver loop (sg_single_driver.cpp:96-124), u and v = (u - u_past)/dt."""
    case = CASES[name]
    got = fullsize.run(Engine, case, case["u0"]())
    compare(name, got, fullsize.oracle_result(name))


def test_config2_sv_persistent_variant_full_size():
    """the Stoermer-Verlet persistent kernel (NLS_FLAG_PK_SV) at 1024^2 x 2000 steps"""
    case = CASES["c2_sv_1024"]
    got = fullsize.run(Engine, case, case["u0"](), flags=nls_b200.FLAG_PK_SV)
    compare("c2_sv_1024[pk]", got, fullsize.oracle_result("c2_sv_1024"))


@pytest.mark.parametrize("name,flags", [("c3_64", 0), ("c3_64", nls_b200.FLAG_NO_PERSISTENT),
                                        ("c3_128", nls_b200.FLAG_NO_PERSISTENT), ("c3_256", 0)])
def test_config3_full_size(name, flags):
    """BASELINE configs[2]: NLSE 3D cubic-quintic, div(c grad) builder semantics, m = 12: 64^3 x 200 steps,
    128^3 x 20 steps through the multi-kernel path (the path 256^3 takes), and 256^3 itself (the bench's
    grid: 296-CTA grids, first-chunk split of pass A) for 2 steps."""
    case = CASES[name]
    got = fullsize.run(Engine, case, case["u0"](), flags=flags)
    compare(f"{name}[flags={flags}]", got, fullsize.oracle_result(name))


def test_config3_256_first_chunk_split_cross_check():
    """256^3: the default pass-A column split vs one handle with NLS_FIRST_CAP=12 (no split), 3 steps"""
    case = dict(CASES["c3_256"], cps=(3,))
    u0 = case["u0"]()
    a = fullsize.run(Engine, case, u0)[3][0]
    os.environ["NLS_FIRST_CAP"] = "12"
    try:
        b = fullsize.run(Engine, case, u0)[3][0]
    finally:
        del os.environ["NLS_FIRST_CAP"]
    err = rel_l2(a, b)
    print(f"c3_256 default split vs NLS_FIRST_CAP=12: {err:.3e}")
    assert err <= 1e-13


def test_config4_batch64_full_size():
    """BASELINE configs[3]: one per-GPU batch of 64 trajectories at 512^2 (the bench's launch shape),
    200 steps; ensemble members t = 0, 17, 255, 511 of the seed-10000+t family against the oracle."""
    case = CASES["c4_512"]
    got = fullsize.run(Engine, case, case["u0"]())
    compare("c4_512 members (0,17,255,511)", got, fullsize.oracle_result("c4_512"), members=case["oracle_members"])


@pytest.mark.parametrize("name", ["c5_128", "c5_1024x16"])
def test_config5_single_gpu_full_size(name):
    """BASELINE configs[4] on one GPU: the slab workload's equation and IC on a 128^3 cube (10 steps) and on a
    slab-shaped 1024 x 1024 x 16 grid (16 MiB planes, 2 steps) against the oracle; the slab-count invariance
    at 256^3 needs 2 GPUs and lives in tests/slab_worker.py."""
    case = CASES[name]
    got = fullsize.run(Engine, case, case["u0"]())
    compare(name, got, fullsize.oracle_result(name))


@pytest.mark.parametrize("m", [20, 25, 30])
def test_persistent_kernel_large_krylov_white_noise(oracle_lib, m):
    """the persistent kernel's one-pass block re-orthogonalisation at m up to 30 with a
    white-noise state (no Krylov convergence to hide behind), against the oracle's sequential MGS sweep."""
    n, L, dt, steps = 96, 5.0, 1e-3, 3
    rng = np.random.default_rng(42)
    u0 = (rng.standard_normal((1, n, n)) + 1j * rng.standard_normal((1, n, n))) / n
    res = {}
    for flags in (0, nls_b200.FLAG_NO_PERSISTENT):
        d = Desc.make(n, n, L=L, equation=EQ.NLSE_CUBIC, krylov_m=m, flags=flags)
        e = Engine(d)
        e.set(FIELD.M, np.ones(d.n))
        e.set(FIELD.U, u0)
        e.step(dt, steps)
        res[flags] = (e.get(), e.launch_count())
    o = oracle_lib.Oracle(d)
    o.set(FIELD.M, np.ones(d.n))
    o.set(FIELD.U, u0)
    o.step(dt, steps)
    ref = o.get()
    e_pk, e_mk = rel_l2(res[0][0], ref), rel_l2(res[nls_b200.FLAG_NO_PERSISTENT][0], ref)
    print(f"white noise m={m}: persistent vs oracle {e_pk:.3e} ({res[0][1]} launches), multi-kernel vs oracle {e_mk:.3e}")
    assert e_pk <= 1e-10 and e_mk <= 1e-10


def test_persistent_kernel_barrier_timeout_is_reported():
    """a grid barrier that times out must poison the handle instead of returning NLS_OK on a
    partially advanced state.  NLS_FLAG_DEBUG_PK_FAULT makes one CTA skip its first barrier (~2 s time-out)."""
    n = 128
    d = Desc.make(n, n, L=10.0, equation=EQ.NLSE_CUBIC, krylov_m=10, flags=nls_b200.FLAG_DEBUG_PK_FAULT)
    e = Engine(d)
    e.set(FIELD.M, np.ones(d.n))
    e.set(FIELD.U, ics.gaussian_bumps_2d(n, 10.0)[None])
    with pytest.raises(nls_b200.NlsError, match="grid barrier"):
        e.step(1e-3, 3)  # nls_step is asynchronous; the nls_sync behind it reports the abort
    with pytest.raises(nls_b200.NlsError, match="poisoned"):
        e.get()
    with pytest.raises(nls_b200.NlsError, match="poisoned"):
        e.step(1e-3, 1)
    # an untouched handle on the same device still works
    d2 = Desc.make(n, n, L=10.0, equation=EQ.NLSE_CUBIC, krylov_m=10)
    e2 = Engine(d2)
    e2.set(FIELD.M, np.ones(d2.n))
    e2.set(FIELD.U, ics.gaussian_bumps_2d(n, 10.0)[None])
    e2.step(1e-3, 3)
    assert np.isfinite(e2.get()).all()


def test_persistent_kernel_long_run_is_split():
    """more steps in one nls_step call than one launch's barrier counter can hold"""
    n = 16
    d = Desc.make(n, n, L=10.0, equation=EQ.SGE, integrator=INT.SV, dt=1e-4, flags=nls_b200.FLAG_PK_SV)
    u0 = ics.kink_2d(n, 10.0)[None]
    e = Engine(d)
    e.set(FIELD.M, np.ones(d.n)); e.set(FIELD.U, u0); e.set(FIELD.V, np.zeros_like(u0))
    e.step(1e-4, 70000)
    assert e.launch_count() >= 3
    e2 = Engine(d)
    e2.set(FIELD.M, np.ones(d.n)); e2.set(FIELD.U, u0); e2.set(FIELD.V, np.zeros_like(u0))
    for k in (30000, 40000):
        e2.step(1e-4, k)
    assert np.array_equal(e.get(), e2.get())


# ---- temporally blocked Stoermer-Verlet kernel (k_pk_sv2: 2D tiles + halo rings in shared memory) ------------
@pytest.mark.parametrize("nx,ny,steps,eq", [(131, 131, 61, EQ.SGE), (16, 16, 9, EQ.SGE), (20, 37, 23, EQ.KGE),
                                            (300, 77, 40, EQ.PHI4), (512, 512, 50, EQ.SGE_DOUBLE),
                                            (257, 640, 33, EQ.SGE_SINH), (1024, 1024, 37, EQ.SGE)])
def test_sv2_tiles_match_oracle_and_graph_path(oracle_lib, nx, ny, steps, eq):
    """default Stoermer-Verlet path for 2D uniform grids: every tile shape (ragged last tiles, tiles on the
    domain boundary, non-square grids), step counts that are not multiples of the block depth, several
    nls_step calls (buffer-role exchange), against the oracle and the one-kernel-per-step graph path."""
    L, dt = 10.0, 1e-3
    big = max(nx, ny)
    u0 = ics.kink_2d(big, L)[:ny, :nx][None]
    if eq in (EQ.KGE, EQ.PHI4, EQ.SGE_SINH):
        u0 = 0.2 * u0
    v0 = 0.05 * np.cos(u0)
    mf = 1.0 + 0.2 * np.random.default_rng(1).standard_normal(nx * ny)
    outs = []
    for flags in (0, nls_b200.FLAG_NO_PERSISTENT):
        d = Desc.make(nx, ny, dx=2 * L / (big - 1), equation=eq, integrator=INT.SV, dt=dt, flags=flags)
        e = Engine(d)
        e.set(FIELD.M, mf); e.set(FIELD.U, u0); e.set(FIELD.V, v0)
        k1 = steps // 3
        for k in (k1, 1, steps - k1 - 1):
            e.step(dt, k)
        outs.append((e.get(), e.get(FIELD.U_PAST), e.get(FIELD.V), e.launch_count()))
    assert outs[0][3] == 3 + 2, "temporally blocked kernel: one launch per nls_step call (+ set V, get V)"
    o = oracle_lib.Oracle(d)
    o.set(FIELD.M, mf); o.set(FIELD.U, u0); o.set(FIELD.V, v0)
    o.step(dt, steps)
    eu, ep = rel_l2(outs[0][0], o.get()), rel_l2(outs[0][1], o.get(FIELD.U_PAST))
    print(f"sv2 {nx}x{ny} x{steps}: vs oracle u {eu:.3e} u_past {ep:.3e}; vs graph path {rel_l2(outs[0][0], outs[1][0]):.3e}")
    assert eu <= 1e-11 and ep <= 1e-11
    assert rel_l2(outs[0][2], o.get(FIELD.V)) <= 1e-7
    assert rel_l2(outs[0][0], outs[1][0]) <= 1e-13


def test_sv2_barrier_timeout_is_reported():
    n = 128
    d = Desc.make(n, n, L=10.0, equation=EQ.SGE, integrator=INT.SV, dt=1e-3, flags=nls_b200.FLAG_DEBUG_PK_FAULT)
    e = Engine(d)
    u0 = ics.kink_2d(n, 10.0)[None]
    e.set(FIELD.M, np.ones(d.n)); e.set(FIELD.U, u0); e.set(FIELD.V, np.zeros_like(u0))
    with pytest.raises(nls_b200.NlsError, match="grid barrier"):
        e.step(1e-3, 40)
    with pytest.raises(nls_b200.NlsError, match="poisoned"):
        e.get()
