"""ctypes binding of the CPU oracle (oracle/nls_oracle.cpp).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))
import nls_b200  # noqa: E402  (descriptor + enums are shared with the product ABI)
from nls_b200 import Desc, FIELD  # noqa: E402

_LIB = os.path.join(_HERE, "_build", "libnls_oracle.so")
_vp, _dp = C.c_void_p, C.POINTER(C.c_double)
_lib = None


def build(force=False):
    if force or not os.path.exists(_LIB) or any(
            os.path.getmtime(os.path.join(_HERE, f)) > os.path.getmtime(_LIB)
            for f in ("nls_oracle.cpp", "nls_oracle.h")):
        subprocess.run(["make", "-C", _HERE, "-B", "_build/libnls_oracle.so"], check=True,
                       stdout=subprocess.DEVNULL)
    return _LIB


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        L = C.CDLL(_LIB)
        L.orc_last_error.restype = C.c_char_p
        L.orc_create.argtypes = [C.POINTER(Desc), C.POINTER(_vp)]
        L.orc_destroy.argtypes = [_vp]
        L.orc_set_field.argtypes = [_vp, C.c_int, _vp, C.c_size_t]
        L.orc_get_field.argtypes = [_vp, C.c_int, _vp, C.c_size_t]
        L.orc_step.argtypes = [_vp, C.c_double, C.c_uint32]
        L.orc_stencil_apply.argtypes = [_vp, _vp, _vp]
        L.orc_matfunc.argtypes = [_vp, C.c_int, C.c_double, C.c_double, _vp, _vp]
        L.orc_lanczos.argtypes = [_vp, _vp, _vp, _vp, _dp]
        L.orc_reduce.argtypes = [_vp, C.c_int, _dp]
        L.orc_get_field_resampled.argtypes = [_vp, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, _vp, C.c_size_t]
        L.orc_csr.argtypes = [_vp, C.POINTER(C.c_int64), _vp, _vp, _vp]
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(_vp)


class Oracle:
    """Same surface as nls_b200.Engine, computed by the CPU restatement."""

    def __init__(self, desc):
        self.lib = lib()
        self.desc = desc
        self.h = _vp()
        self._check(self.lib.orc_create(C.byref(desc), C.byref(self.h)))

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError(f"oracle error {rc}: {self.lib.orc_last_error().decode()}")

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.orc_destroy(self.h)
            self.h = _vp()

    def _state_shape(self):
        d = self.desc
        return (d.batch, d.nz, d.ny, d.nx) if d.nz > 1 else (d.batch, d.ny, d.nx)

    def set(self, field, arr):
        dt = self.desc.dtype if field in (FIELD.U, FIELD.U_PAST, FIELD.V) else np.float64
        a = np.ascontiguousarray(arr, dtype=dt)
        self._check(self.lib.orc_set_field(self.h, field, _ptr(a), a.nbytes))

    def get(self, field=FIELD.U):
        out = np.empty(self._state_shape(), dtype=self.desc.dtype)
        self._check(self.lib.orc_get_field(self.h, field, _ptr(out), out.nbytes))
        return out

    def get_resampled(self, tnx, tny, tnz=1, field=FIELD.U):
        shape = (self.desc.batch, tnz, tny, tnx) if self.desc.nz > 1 else (self.desc.batch, tny, tnx)
        out = np.empty(shape, dtype=self.desc.dtype)
        self._check(self.lib.orc_get_field_resampled(self.h, field, tnx, tny, tnz, _ptr(out), out.nbytes))
        return out

    def step(self, dt, nsteps=1, sync=True):
        self._check(self.lib.orc_step(self.h, dt, nsteps))

    def stencil_apply(self, x):
        x = np.ascontiguousarray(x, dtype=self.desc.dtype)
        y = np.empty_like(x)
        self._check(self.lib.orc_stencil_apply(self.h, _ptr(x), _ptr(y)))
        return y

    def matfunc(self, fn, t, x):
        x = np.ascontiguousarray(x, dtype=self.desc.dtype)
        y = np.empty_like(x)
        t = complex(t)
        self._check(self.lib.orc_matfunc(self.h, fn, t.real, t.imag, _ptr(x), _ptr(y)))
        return y

    def lanczos(self, x):
        d = self.desc
        x = np.ascontiguousarray(x, dtype=d.dtype)
        V = np.empty((d.krylov_m, d.n), dtype=d.dtype)
        alpha = np.empty(d.krylov_m, dtype=d.dtype)
        beta = np.empty(d.krylov_m, dtype=np.float64)
        self._check(self.lib.orc_lanczos(self.h, _ptr(x), _ptr(V), _ptr(alpha), beta.ctypes.data_as(_dp)))
        return V, alpha, beta

    def reduce(self, what):
        out = np.empty(self.desc.batch, dtype=np.float64)
        self._check(self.lib.orc_reduce(self.h, what, out.ctypes.data_as(_dp)))
        return out

    def csr(self):
        import scipy.sparse as sp
        nnz = C.c_int64()
        self._check(self.lib.orc_csr(self.h, C.byref(nnz), None, None, None))
        n = self.desc.n
        rp = np.empty(n + 1, np.int64)
        col = np.empty(nnz.value, np.int32)
        val = np.empty(nnz.value, np.float64)
        self._check(self.lib.orc_csr(self.h, C.byref(nnz), _ptr(rp), _ptr(col), _ptr(val)))
        return sp.csr_matrix((val, col, rp), shape=(n, n))

    @staticmethod
    def threads():
        return lib().orc_threads()

    @staticmethod
    def set_threads(n):
        lib().orc_set_threads(int(n))
