"""ctypes binding of oracle/_ref/libnls_ref.so: the reference's own host sources
(/root/reference/nlsolvers/{common,host}/include/*.hpp, compiled unmodified against the mini-Eigen
stand-in of oracle/ref_shim by oracle/Makefile).  TEST INFRASTRUCTURE ONLY.  The library can only be
BUILT where /root/reference exists (this container); the built file travels to the GPU box."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_ref", "libnls_ref.so")
REFERENCE = "/root/reference/nlsolvers"
_vp = C.c_void_p
_lib = None


def available():
    return os.path.exists(_LIB) or os.path.isdir(REFERENCE)


def build(force=False):
    """compile the reference headers where they lie; outputs only into oracle/_ref/"""
    if not os.path.isdir(REFERENCE):
        return _LIB if os.path.exists(_LIB) else None
    srcs = [os.path.join(_HERE, "ref_harness.cpp"), os.path.join(_HERE, "ref_shim", "Eigen", "mini_eigen.hpp")]
    if force or not os.path.exists(_LIB) or any(os.path.getmtime(s) > os.path.getmtime(_LIB) for s in srcs):
        subprocess.run(["make", "-C", _HERE, "-B", "_ref/libnls_ref.so"], check=True, stdout=subprocess.DEVNULL)
    return _LIB


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        L = C.CDLL(_LIB)
        L.ref_create.restype = _vp
        L.ref_create.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_double, C.c_double, C.c_double, _vp,
                                 C.c_double]
        L.ref_destroy.argtypes = [_vp]
        L.ref_nnz.restype = C.c_long
        L.ref_nnz.argtypes = [_vp]
        L.ref_matrix.argtypes = [_vp, _vp, _vp, _vp]
        L.ref_apply.argtypes = [_vp, C.c_int, _vp, _vp]
        L.ref_lanczos.argtypes = [_vp, C.c_int, _vp, C.c_uint32, _vp, _vp, C.POINTER(C.c_double)]
        L.ref_matfunc.argtypes = [_vp, C.c_int, C.c_double, C.c_double, _vp, _vp, C.c_uint32]
        L.ref_nlse_steps.argtypes = [_vp, C.c_int, _vp, _vp, _vp, C.c_double, C.c_double, C.c_double, C.c_double,
                                     C.c_uint32]
        L.ref_real_steps.argtypes = [_vp, C.c_int, _vp, _vp, _vp, C.c_double, C.c_uint32]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(_vp)


class Reference:
    """One assembled operator of the reference + its solver functions."""

    def __init__(self, nx, ny, nz=1, dx=1.0, aniso=False, c=None, sign=1.0):
        self.L = lib()
        self.n, self.shape = nx * ny * nz, ((nz, ny, nx) if nz > 1 else (ny, nx))
        kind = (2 if nz > 1 else 0) + (1 if aniso else 0)
        cc = None if c is None else np.ascontiguousarray(c, dtype=np.float64)
        self.h = self.L.ref_create(kind, nx, ny, nz, dx, dx, dx, None if cc is None else _p(cc), sign)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_destroy(self.h)
            self.h = None

    def matrix(self):
        import scipy.sparse as sp
        nnz = self.L.ref_nnz(self.h)
        outer, inner, val = np.empty(self.n + 1, np.int64), np.empty(nnz, np.int32), np.empty(nnz, np.float64)
        self.L.ref_matrix(self.h, _p(outer), _p(inner), _p(val))
        return sp.csc_matrix((val, inner, outer), shape=(self.n, self.n)).tocsr()

    def apply(self, x):
        x = np.ascontiguousarray(x)
        y = np.empty_like(x)
        self.L.ref_apply(self.h, int(np.iscomplexobj(x)), _p(x), _p(y))
        return y

    def lanczos(self, u, m):
        u = np.ascontiguousarray(u).ravel()
        cp = np.iscomplexobj(u)
        V = np.empty((m, self.n), dtype=u.dtype)  # column major n x m == C-order (m, n)
        T = np.empty((m, m), dtype=u.dtype)
        beta = C.c_double()
        self.L.ref_lanczos(self.h, int(cp), _p(u), m, _p(V), _p(T), C.byref(beta))
        return V, T.T.copy(), beta.value

    def matfunc(self, fn, t, x, m=10):
        x = np.ascontiguousarray(x, dtype=np.complex128 if fn <= 1 else np.float64)
        y = np.empty_like(x)
        t = complex(t)
        self.L.ref_matfunc(self.h, fn, t.real, t.imag, _p(x), _p(y), m)
        return y

    def nlse_steps(self, kind, u, mfield, dt, nsteps, uprev=None, s1=0.0, s2=0.0, kappa=0.0):
        u = np.ascontiguousarray(u, dtype=np.complex128).copy()
        up = None if uprev is None else np.ascontiguousarray(uprev, dtype=np.complex128).copy()
        mf = np.ascontiguousarray(mfield, dtype=np.float64)
        self.L.ref_nlse_steps(self.h, kind, _p(u), None if up is None else _p(up), _p(mf), dt, s1, s2, kappa, nsteps)
        return (u, up) if up is not None else u

    def real_steps(self, kind, u, upast, mfield, dt, nsteps):
        u = np.ascontiguousarray(u, dtype=np.float64).copy()
        up = np.ascontiguousarray(upast, dtype=np.float64).copy()
        mf = np.ascontiguousarray(mfield, dtype=np.float64)
        self.L.ref_real_steps(self.h, kind, _p(u), _p(up), _p(mf), dt, nsteps)
        return u, up
