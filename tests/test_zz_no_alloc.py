"""nls_reduce_mass_noalloc (include/nls.h): the allocation-free variant of the NLS_RED_MASS reduction, whose grid-
reduction scratch is a file-scope static __device__ array instead of the handle's cudaMalloc'ed buffers.  It has no
counterpart in the reference; the checker is the existing entry point,
nls_reduce(h, NLS_RED_MASS), on the same state, rtol 1e-12 like tests/test_energy.py.

Collected last on purpose (file name).  The inputs / outputs of the new entry point are
the caller's device arrays -- torch tensors here, allocated by the TEST, not by the entry point."""
import ctypes as C

import numpy as np
import pytest

import ics
import nls_b200
from nls_b200 import Desc, Engine, EQ, INT, FIELD, RED

pytestmark = pytest.mark.gpu

RTOL = 1e-12
NLS_ERR_ARG = -1


def _mass_existing(d, u0):
    e = Engine(d)
    e.set(FIELD.M, np.ones(d.n))
    e.set(FIELD.U, u0)
    return e.reduce(RED.MASS)


def _mass_noalloc(lib, u0, batch, cell, stride=0):
    import torch
    cplx = np.iscomplexobj(u0)
    host = np.ascontiguousarray(u0, dtype=np.complex128 if cplx else np.float64).reshape(batch, -1)
    n = host.shape[1]
    if stride:  # trajectories `stride` elements apart, NaN in the gaps: they must not be read
        padded = np.full((batch, stride), np.nan, dtype=host.dtype)
        padded[:, :n] = host
        host = padded
    dev_u = torch.from_numpy(host).cuda()
    dev_out = torch.full((batch,), -1.0, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.nls_reduce_mass_noalloc(0, C.c_void_p(dev_u.data_ptr()), n, batch, stride, int(cplx), cell,
                                     C.c_void_p(dev_out.data_ptr()), C.c_void_p(stream))
    assert rc == 0, lib.nls_last_error().decode()
    torch.cuda.synchronize()
    return dev_out.cpu().numpy()


def test_noalloc_mass_matches_nls_reduce_on_two_calls_in_a_row():
    lib = nls_b200.load_library()
    n, L = 96, 10.0
    # call 1: complex state, 3 trajectories
    d1 = Desc.make(n, n, L=L, equation=EQ.NLSE_CUBIC, krylov_m=10, batch=3)
    u1 = np.stack([ics.gaussian_bumps_2d(n, L, seed=10 + b) for b in range(3)])
    want1 = _mass_existing(d1, u1)
    got1 = _mass_noalloc(lib, u1, 3, d1.dx * d1.dy)
    # call 2 straight after, different shape and type: the shared scratch still holds call 1's partial sums
    d2 = Desc.make(64, 48, L=L, equation=EQ.SGE, integrator=INT.GAUTSCHI, krylov_m=10, dt=1e-3, batch=1)
    u2 = np.random.default_rng(5).standard_normal(d2.n)
    want2 = _mass_existing(d2, u2)
    got2 = _mass_noalloc(lib, u2, 1, d2.dx * d2.dy)
    # call 3: call 1 again with padded trajectory stride -> same numbers
    got3 = _mass_noalloc(lib, u1, 3, d1.dx * d1.dy, stride=n * n + 37)
    print("no-alloc mass: worst rel err", max(np.max(np.abs(got1 - want1) / want1), np.max(np.abs(got2 - want2) / want2)))
    assert np.allclose(got1, want1, rtol=RTOL, atol=0), (got1, want1)
    assert np.allclose(got2, want2, rtol=RTOL, atol=0), (got2, want2)
    assert np.allclose(got3, want1, rtol=RTOL, atol=0), (got3, want1)
    assert np.all(want1 > 0) and want2[0] > 0


def test_noalloc_mass_large_single_trajectory():
    """more CTAs per trajectory than one wave: the partial-slot cap of the static scratch is exercised"""
    lib = nls_b200.load_library()
    n = 1024
    d = Desc.make(n, n, L=10.0, equation=EQ.NLSE_CUBIC, krylov_m=10)
    rng = np.random.default_rng(11)
    u = rng.standard_normal(n * n) + 1j * rng.standard_normal(n * n)
    want = _mass_existing(d, u[None])
    for _ in range(2):
        got = _mass_noalloc(lib, u[None], 1, d.dx * d.dy)
        assert np.allclose(got, want, rtol=RTOL, atol=0), (got, want)
    assert want[0] == pytest.approx(np.sum(np.abs(u) ** 2) * d.dx * d.dy, rel=1e-12)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    lib = nls_b200.load_library()
    import torch
    buf = torch.zeros(1024, dtype=torch.float64, device="cuda")
    p = C.c_void_p(buf.data_ptr())
    rc = lib.nls_reduce_mass_noalloc(0, p, 4, 257, 0, 0, 1.0, p, None)  # batch > NLS_NOALLOC_MAX_BATCH
    assert rc == NLS_ERR_ARG and b"NLS_NOALLOC_MAX_BATCH" in lib.nls_last_error()
    rc = lib.nls_reduce_mass_noalloc(0, p, (1 << 31) + 1, 1, 0, 0, 1.0, p, None)  # n_sites > NLS_NOALLOC_MAX_SITES
    assert rc == NLS_ERR_ARG and b"NLS_NOALLOC_MAX_SITES" in lib.nls_last_error()
    rc = lib.nls_reduce_mass_noalloc(0, None, 4, 1, 0, 0, 1.0, p, None)
    assert rc == NLS_ERR_ARG
