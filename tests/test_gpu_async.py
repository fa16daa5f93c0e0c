"""Asynchronous snapshot egress / state upload (nls_get_field_async, nls_get_field_resampled_async,
nls_set_field_async, nls_wait): same bytes as the blocking calls, taken at the right point of
the handle's stream order while later steps are already queued."""
import numpy as np
import pytest

import ics
import nls_b200
from nls_b200 import Desc, Engine, EQ, INT, FIELD

pytestmark = pytest.mark.gpu


def _mk(kind, batch=1, n=96):
    L, dt = 10.0, 1e-3
    if kind == "nlse":
        d = Desc.make(n, n, L=L, equation=EQ.NLSE_CUBIC, krylov_m=10, batch=batch)
        u0 = np.stack([ics.gaussian_bumps_2d(n, L, seed=10 + b) for b in range(batch)])
        v0 = None
    elif kind == "sewi":
        d = Desc.make(n, n, L=L, equation=EQ.NLSE_CUBIC, integrator=INT.SEWI, krylov_m=8, batch=batch)
        u0 = np.stack([ics.gaussian_bumps_2d(n, L, seed=10 + b) for b in range(batch)])
        v0 = None
    else:
        d = Desc.make(n, n, L=L, equation=EQ.SGE, integrator=INT.GAUTSCHI if kind == "sge" else INT.SV, krylov_m=10, dt=dt,
                      batch=batch)
        u0 = np.stack([ics.kink_2d(n, L, seed=70 + b) for b in range(batch)])
        v0 = 0.05 * np.cos(u0)
    return d, u0, v0, dt


def _load(e, d, u0, v0):
    e.set(FIELD.M, np.ones(d.n))
    e.set(FIELD.U, u0)
    if v0 is not None:
        e.set(FIELD.V, v0)


@pytest.mark.parametrize("kind,batch", [("nlse", 1), ("nlse", 3), ("sge", 1), ("sge_sv", 1), ("sge_sv", 2), ("sewi", 1)])
def test_async_snapshots_equal_blocking_snapshots(kind, batch):
    d, u0, v0, dt = _mk(kind, batch)
    ea, eb = Engine(d), Engine(d)
    _load(ea, d, u0, v0)
    _load(eb, d, u0, v0)
    fields = [FIELD.U] + ([FIELD.V, FIELD.U_PAST] if v0 is not None else []) + ([FIELD.U_PAST] if kind == "sewi" else [])
    nsnap, seg = 5, 4
    shape = (nsnap, batch) + u0.shape[1:]
    pins = {f: nls_b200.pinned_empty(shape, d.dtype) for f in fields}
    ref = {f: np.empty(shape, d.dtype) for f in fields}
    for s in range(nsnap):  # everything queued without waiting: 5 segments, 5 x len(fields) snapshots in flight
        ea.step(dt, seg, sync=False)
        for f in fields:
            ea.get_async(f, pins[f][s])
    ea.wait()
    for s in range(nsnap):
        eb.step(dt, seg)
        for f in fields:
            ref[f][s] = eb.get(f)
    for f in fields:
        assert np.array_equal(pins[f], ref[f]), f"field {f}"
    for a in pins.values():
        nls_b200.pinned_free(a)


def test_async_batch_lands_in_trajectory_layout():
    """dst_traj_stride: an ensemble's snapshot goes straight into a (B, num_snapshots, ny, nx) file image"""
    d, u0, v0, dt = _mk("nlse", batch=4, n=64)
    e, eb = Engine(d), Engine(d)
    _load(e, d, u0, None)
    _load(eb, d, u0, None)
    ns = 3
    traj = nls_b200.pinned_empty((4, ns, 64, 64), np.complex128)
    traj[...] = 0
    stride = ns * 64 * 64 * 16
    for s in range(ns):
        e.step(dt, 2, sync=False)
        e.get_async(FIELD.U, traj[0, s:], traj_stride=stride)
    e.wait()
    for s in range(ns):
        eb.step(dt, 2)
        assert np.array_equal(traj[:, s], eb.get())
    nls_b200.pinned_free(traj)


@pytest.mark.parametrize("kind", ["nlse", "sge"])
def test_async_resampled_equals_blocking(kind):
    d, u0, v0, dt = _mk(kind, batch=2, n=80)
    e, eb = Engine(d), Engine(d)
    _load(e, d, u0, v0)
    _load(eb, d, u0, v0)
    fields = [FIELD.U] + ([FIELD.V] if v0 is not None else [])
    out = {f: nls_b200.pinned_empty((3, 2, 32, 24), d.dtype) for f in fields}
    for s in range(3):
        e.step(dt, 3, sync=False)
        for f in fields:
            e.get_resampled_async(f, 24, 32, 1, out[f][s])
    e.wait()
    for s in range(3):
        eb.step(dt, 3)
        for f in fields:
            assert np.array_equal(out[f][s], eb.get_resampled(24, 32, 1, field=f))
    for a in out.values():
        nls_b200.pinned_free(a)


@pytest.mark.parametrize("kind", ["nlse", "sge_sv", "sewi"])
def test_async_upload_equals_blocking_upload(kind):
    d, u0, v0, dt = _mk(kind, batch=2, n=72)
    ea, eb = Engine(d), Engine(d)
    _load(ea, d, u0, v0)
    _load(eb, d, u0, v0)
    ea.step(dt, 3, sync=False)
    eb.step(dt, 3)
    # restart both from a different state while the first handle's steps may still be running
    u1 = nls_b200.pinned_empty(u0.shape, d.dtype)
    u1[...] = u0[::-1] * 0.9
    ea.set_async(FIELD.U, u1)
    eb.set(FIELD.U, np.array(u1))
    if v0 is not None:
        p1 = nls_b200.pinned_empty(u0.shape, d.dtype)
        p1[...] = u1 - dt * v0
        ea.set_async(FIELD.U_PAST, p1)
        eb.set(FIELD.U_PAST, np.array(p1))
    ea.step(dt, 4, sync=False)
    eb.step(dt, 4)
    ea.wait()
    assert np.array_equal(ea.get(), eb.get())
    nls_b200.pinned_free(u1)


def test_pinned_source_set_field_does_not_need_a_sync():
    """a page-locked source makes nls_set_field asynchronous; results are identical"""
    d, u0, v0, dt = _mk("nlse", batch=1, n=64)
    ea, eb = Engine(d), Engine(d)
    pu = nls_b200.pinned_empty(u0.shape, d.dtype)
    pu[...] = u0
    ea.set(FIELD.M, np.ones(d.n)); eb.set(FIELD.M, np.ones(d.n))
    ea.set(FIELD.U, pu)
    eb.set(FIELD.U, u0)
    ea.step(dt, 5); eb.step(dt, 5)
    assert np.array_equal(ea.get(), eb.get())
    nls_b200.pinned_free(pu)
