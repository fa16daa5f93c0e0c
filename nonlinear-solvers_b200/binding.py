"""ctypes mirror of include/nls.h (the C ABI of libnls_b200.so).

No torch types cross this boundary: host numpy arrays or raw device pointers only.
There is NO CPU fallback: if the CUDA library is missing, or no GPU is visible,
every compute call raises NlsError.
"""
import ctypes as C
import os
import subprocess
import sys
from types import SimpleNamespace

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
_CSRC = os.path.join(_PKG, "csrc")
_LIB = os.path.join(_PKG, "libnls_b200.so")

EQ = SimpleNamespace(NLSE_CUBIC=0, NLSE_CQ=1, NLSE_SAT=2, SGE=3, KGE=4, PHI4=5, SGE_DOUBLE=6, SGE_SINH=7)
INT = SimpleNamespace(SS2=0, GAUTSCHI=1, SV=2, SEWI=3)
LAP = SimpleNamespace(UNIFORM=0, ANISO=1)
REORTH = SimpleNamespace(FULL=0, NONE=1, MGS=2, GRAM=3, BLOCK=4)
FN = SimpleNamespace(EXP=0, SINC=1, COS_SQRT=2, SINC2_SQRT=3, ID_SQRT=4, SINC2_HALF=5, MOD_COSINE=6)
FIELD = SimpleNamespace(U=0, U_PAST=1, V=2, M=3, C=4)
RED = SimpleNamespace(MASS=0, NONFINITE=1, ENERGY=2)
FLAG_NO_GRAPH = 1
FLAG_NO_PERSISTENT = 2
FLAG_SEWI_HOST2D = 4
FLAG_MARCH = 8
FLAG_TMA = 32
FLAG_PK_SV = 64
FLAG_TMAP = 128
FLAG_NO_TMAP = 256
FLAG_DEBUG_PK_FAULT = 0x80000000


class NlsError(RuntimeError):
    pass


class GenTerm(C.Structure):
    """struct nls_gen_term (include/nls.h): one term of an on-device field generator."""
    _fields_ = [("kind", C.c_int32), ("iparam", C.c_int32), ("p", C.c_double * 10)]


GEN = SimpleNamespace(CONST=0, GAUSS=1, COSWAVE=2, TANH_BOUNDARY=3, SINSIN=4, LAYERS=5, CHECKER=6, HALFSPACE=7,
                      GAUSS_WAVE=8, KINK=9)
GENPOST = SimpleNamespace(NONE=0, CLAMP_MIN=1, MINMAX=2, MAXABS=3, GRADIENT_MASS=4)


class Desc(C.Structure):
    """struct nls_desc (include/nls.h)."""
    _fields_ = [
        ("struct_size", C.c_uint32), ("nx", C.c_uint32), ("ny", C.c_uint32), ("nz", C.c_uint32),
        ("dx", C.c_double), ("dy", C.c_double), ("dz", C.c_double),
        ("equation", C.c_int32), ("integrator", C.c_int32), ("laplacian", C.c_int32),
        ("reorth", C.c_int32), ("krylov_m", C.c_uint32), ("batch", C.c_uint32),
        ("s1", C.c_double), ("s2", C.c_double), ("kappa", C.c_double),
        ("op_sign", C.c_double), ("dt", C.c_double), ("device", C.c_int32),
        ("nz_global", C.c_uint32), ("z0", C.c_uint32), ("flags", C.c_uint32),
    ]

    @classmethod
    def make(cls, nx, ny, nz=1, L=None, dx=None, equation=0, integrator=0, laplacian=0, reorth=0,
             krylov_m=10, batch=1, s1=0.0, s2=0.0, kappa=0.0, op_sign=1.0, dt=0.0, device=0,
             nz_global=0, z0=0, flags=0):
        """dx follows the drivers: dx = 2*Lx/(nx-1) (nlse_cubic_driver.cpp:46)."""
        d = cls()
        d.struct_size = C.sizeof(cls)
        d.nx, d.ny, d.nz = nx, ny, nz
        if dx is None:
            dx = 2.0 * L / (nx - 1)
        d.dx = d.dy = d.dz = dx
        d.equation, d.integrator, d.laplacian, d.reorth = equation, integrator, laplacian, reorth
        d.krylov_m, d.batch = krylov_m, batch
        d.s1, d.s2, d.kappa, d.op_sign, d.dt, d.device = s1, s2, kappa, op_sign, dt, device
        d.nz_global, d.z0, d.flags = nz_global, z0, flags
        return d

    @property
    def n(self):
        return self.nx * self.ny * self.nz

    @property
    def is_complex(self):
        return self.equation <= EQ.NLSE_SAT

    @property
    def dtype(self):
        return np.complex128 if self.is_complex else np.float64


NVCC_FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC", "-shared", "--expt-relaxed-constexpr"]


def library_path():
    return _LIB


def build_library(force=False, verbose=False):
    """Compile csrc/*.cu for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    srcs = sorted(os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith(".cu"))
    deps = srcs + [os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith((".cuh", ".h"))]
    deps.append(os.path.join(_ROOT, "include", "nls.h"))
    if not force and os.path.exists(_LIB) and all(os.path.getmtime(_LIB) >= os.path.getmtime(s) for s in deps):
        return _LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + ["-I", os.path.join(_ROOT, "include"), "-o", _LIB] + srcs + ["-lcudart"]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd), file=sys.stderr)
    subprocess.run(cmd, check=True)
    return _LIB


_lib = None
_vp, _sz, _dp = C.c_void_p, C.c_size_t, C.POINTER(C.c_double)
_SIGS = {
    "nls_last_error": (C.c_char_p, []),
    "nls_version": (C.c_int, []),
    "nls_device_count": (C.c_int, []),
    "nls_create": (C.c_int, [C.POINTER(Desc), C.POINTER(_vp)]),
    "nls_destroy": (C.c_int, [_vp]),
    "nls_set_field": (C.c_int, [_vp, C.c_int, _vp, _sz]),
    "nls_get_field": (C.c_int, [_vp, C.c_int, _vp, _sz]),
    "nls_get_field_resampled": (C.c_int, [_vp, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, _vp, _sz]),
    "nls_set_field_dev": (C.c_int, [_vp, C.c_int, _vp, _sz]),
    "nls_get_field_dev": (C.c_int, [_vp, C.c_int, _vp, _sz]),
    "nls_step": (C.c_int, [_vp, C.c_double, C.c_uint32]),
    "nls_sync": (C.c_int, [_vp]),
    "nls_stencil_apply": (C.c_int, [_vp, _vp, _vp]),
    "nls_matfunc": (C.c_int, [_vp, C.c_int, C.c_double, C.c_double, _vp, _vp]),
    "nls_lanczos": (C.c_int, [_vp, _vp, _vp, _vp, _dp]),
    "nls_reduce": (C.c_int, [_vp, C.c_int, _dp]),
    "nls_launch_count": (C.c_uint64, [_vp]),
    "nls_tma_launch_count": (C.c_uint64, [_vp]),
    "nls_tmap_launch_count": (C.c_uint64, [_vp]),
    "nls_stream": (_vp, [_vp]),
    "nls_step_timed": (C.c_int, [_vp, C.c_double, C.c_uint32, C.POINTER(C.c_float)]),
    "nls_get_field_async": (C.c_int, [_vp, C.c_int, _vp, _sz, _sz]),
    "nls_get_field_resampled_async": (C.c_int, [_vp, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, _vp, _sz, _sz]),
    "nls_set_field_async": (C.c_int, [_vp, C.c_int, _vp, _sz]),
    "nls_wait": (C.c_int, [_vp]),
    "nls_host_alloc": (C.c_int, [C.POINTER(_vp), _sz]),
    "nls_host_free": (C.c_int, [_vp]),
    "nls_host_register": (C.c_int, [_vp, _sz]),
    "nls_host_unregister": (C.c_int, [_vp]),
    "nls_stencil_apply_dev": (C.c_int, [_vp, _vp, _vp]),
    "nls_matfunc_dev": (C.c_int, [_vp, C.c_int, C.c_double, C.c_double, _vp, _vp]),
    "nls_csr_residual": (C.c_int, [_vp, _vp, _vp, _vp, C.c_uint32, C.c_uint64, C.c_int, C.c_int, _dp]),
    "nls_pointer_is_device": (C.c_int, [_vp]),
    "nls_memcpy_to_host": (C.c_int, [_vp, _vp, _sz]),
    "nls_reduce_mass_noalloc": (C.c_int, [C.c_int, _vp, C.c_uint64, C.c_uint32, C.c_uint64, C.c_int, C.c_double, _vp, _vp]),
    "nls_generate_field": (C.c_int, [_vp, C.c_int, C.c_int32, _vp, C.c_uint32, C.c_int, _dp, C.c_double, C.c_double, C.c_double]),
    "nls_get_field_fft_resampled": (C.c_int, [_vp, C.c_int, C.c_uint32, C.c_uint32, _vp, _sz]),
    "nls_get_field_fft_resampled_async": (C.c_int, [_vp, C.c_int, C.c_uint32, C.c_uint32, _vp, _sz, _sz]),
    "nls_slab_export": (C.c_int, [_vp, _vp]),
    "nls_slab_connect": (C.c_int, [_vp, _vp, C.c_int, C.c_int]),
}
EXPORTS = tuple(_SIGS)


def load_library(path=None):
    """dlopen libnls_b200.so and declare every entry point of include/nls.h."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("NLS_B200_LIB") or _LIB  # the env override is for A/B builds of the same source tree
    if not os.path.exists(p):
        raise NlsError(f"{p} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(there is no CPU fallback)")
    lib = C.CDLL(p)
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype, fn.argtypes = res, args
    if path is None:
        _lib = lib
    return lib


def _ptr(a):
    return a.ctypes.data_as(_vp)


def pinned_empty(shape, dtype):
    """page-locked numpy array (nls_host_alloc); keep the returned array alive, free with pinned_free()."""
    lib = load_library()
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = _vp()
    rc = lib.nls_host_alloc(C.byref(p), max(n, 1))
    if rc != 0:
        raise NlsError(f"nls error {rc}: {lib.nls_last_error().decode()}")
    buf = (C.c_char * max(n, 1)).from_address(p.value)
    a = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    a.flags.writeable = True
    _pinned[a.ctypes.data] = p
    return a


def pinned_free(a):
    p = _pinned.pop(a.ctypes.data, None)
    if p is not None:
        load_library().nls_host_free(p)


_pinned = {}


class Engine:
    """Host-side handle: mirrors the reference solver classes' life cycle
    (ctor uploads state, step(), transfer back; device/nlse_solver_dev.hpp:53-118)."""

    def __init__(self, desc, lib=None):
        self.lib = lib or load_library()
        self.desc = desc
        self.h = _vp()
        self._check(self.lib.nls_create(C.byref(desc), C.byref(self.h)))

    def _check(self, rc):
        if rc != 0:
            raise NlsError(f"nls error {rc}: {self.lib.nls_last_error().decode()}")

    def close(self):
        if self.h:
            self.lib.nls_destroy(self.h)
            self.h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _state_shape(self):
        d = self.desc
        return (d.batch, d.nz, d.ny, d.nx) if d.nz > 1 else (d.batch, d.ny, d.nx)

    def set(self, field, arr):
        dt = self.desc.dtype if field in (FIELD.U, FIELD.U_PAST, FIELD.V) else np.float64
        a = np.ascontiguousarray(arr, dtype=dt)
        self._check(self.lib.nls_set_field(self.h, field, _ptr(a), a.nbytes))

    def get(self, field=FIELD.U):
        out = np.empty(self._state_shape(), dtype=self.desc.dtype)
        self._check(self.lib.nls_get_field(self.h, field, _ptr(out), out.nbytes))
        return out

    def get_resampled(self, tnx, tny, tnz=1, field=FIELD.U):
        """field resampled on the device to (batch, [tnz,] tny, tnx) (downsampling.py:68-104)."""
        shape = (self.desc.batch, tnz, tny, tnx) if self.desc.nz > 1 else (self.desc.batch, tny, tnx)
        out = np.empty(shape, dtype=self.desc.dtype)
        self._check(self.lib.nls_get_field_resampled(self.h, field, tnx, tny, tnz, _ptr(out), out.nbytes))
        return out

    def get_fft_resampled(self, tnx, tny, field=FIELD.U):
        """field downsampled spectrally on the device to (batch, tny, tnx) (downsampling.py:10-34 downsample_fft)."""
        out = np.empty((self.desc.batch, tny, tnx), dtype=self.desc.dtype)
        self._check(self.lib.nls_get_field_fft_resampled(self.h, field, tnx, tny, _ptr(out), out.nbytes))
        return out

    def get_into(self, field, out):
        """device -> caller-provided host array (e.g. pinned memory)."""
        assert out.flags.c_contiguous and out.dtype == self.desc.dtype
        self._check(self.lib.nls_get_field(self.h, field, _ptr(out), out.nbytes))
        return out

    def get_async(self, field, out, traj_stride=0):
        """snapshot -> caller-provided (page-locked) array on the D2H stream; valid after wait()."""
        assert out.flags.c_contiguous
        nbytes = self.desc.batch * self.desc.n * np.dtype(self.desc.dtype).itemsize
        self._check(self.lib.nls_get_field_async(self.h, field, _ptr(out), nbytes, traj_stride))

    def get_resampled_async(self, field, tnx, tny, tnz, out, traj_stride=0):
        nbytes = self.desc.batch * tnx * tny * tnz * np.dtype(self.desc.dtype).itemsize
        self._check(self.lib.nls_get_field_resampled_async(self.h, field, tnx, tny, tnz, _ptr(out), nbytes, traj_stride))

    def set_async(self, field, arr):
        """upload through the H2D staging ring; `arr` must stay unchanged until wait()."""
        assert arr.flags.c_contiguous
        self._check(self.lib.nls_set_field_async(self.h, field, _ptr(arr), arr.nbytes))

    def wait(self):
        self._check(self.lib.nls_wait(self.h))

    def step(self, dt, nsteps=1, sync=True):
        self._check(self.lib.nls_step(self.h, dt, nsteps))
        if sync:
            self._check(self.lib.nls_sync(self.h))

    def step_timed(self, dt, nsteps):
        ms = C.c_float()
        self._check(self.lib.nls_step_timed(self.h, dt, nsteps, C.byref(ms)))
        return ms.value

    def sync(self):
        self._check(self.lib.nls_sync(self.h))

    def stencil_apply(self, x):
        x = np.ascontiguousarray(x, dtype=self.desc.dtype)
        y = np.empty_like(x)
        self._check(self.lib.nls_stencil_apply(self.h, _ptr(x), _ptr(y)))
        return y

    def matfunc(self, fn, t, x):
        x = np.ascontiguousarray(x, dtype=self.desc.dtype)
        y = np.empty_like(x)
        t = complex(t)
        self._check(self.lib.nls_matfunc(self.h, fn, t.real, t.imag, _ptr(x), _ptr(y)))
        return y

    def lanczos(self, x):
        d = self.desc
        x = np.ascontiguousarray(x, dtype=d.dtype)
        V = np.empty((d.krylov_m, d.n), dtype=d.dtype)
        alpha = np.empty(d.krylov_m, dtype=d.dtype)
        beta = np.empty(d.krylov_m, dtype=np.float64)
        self._check(self.lib.nls_lanczos(self.h, _ptr(x), _ptr(V), _ptr(alpha),
                                         beta.ctypes.data_as(_dp)))
        return V, alpha, beta

    def generate(self, field, terms, post=0, post_params=(0.0, 0.0, 0.0), L=None, traj=-1):
        """fill a field on the device from a term table (nls_generate_field).  terms: [(kind, iparam, [p...]), ...];
        L: box half-length(s), scalar or (Lx, Ly, Lz)."""
        arr = (GenTerm * max(len(terms), 1))()
        for t, (kind, ip, ps) in zip(arr, terms):
            t.kind, t.iparam = int(kind), int(ip)
            for q, v in enumerate(ps):
                t.p[q] = float(v)
        Ls = (1.0, 1.0, 1.0) if L is None else (L, L, L) if np.isscalar(L) else tuple(L) + (L[-1],) * (3 - len(L))
        pp = (C.c_double * 3)(*[float(x) for x in (list(post_params) + [0.0, 0.0, 0.0])[:3]])
        self._check(self.lib.nls_generate_field(self.h, field, traj, arr if terms else None, len(terms), post, pp,
                                                float(Ls[0]), float(Ls[1]), float(Ls[2])))

    def csr_residual(self, csr):
        """max |CSR x - L x| / max |L x| for a scipy.sparse CSR matrix (host arrays) -- nls_csr_residual"""
        rp = np.ascontiguousarray(csr.indptr, dtype=np.int32)
        ci = np.ascontiguousarray(csr.indices, dtype=np.int32)
        cplx = np.iscomplexobj(csr.data)
        val = np.ascontiguousarray(csr.data, dtype=np.complex128 if cplx else np.float64)
        err = C.c_double()
        self._check(self.lib.nls_csr_residual(self.h, _ptr(rp), _ptr(ci), _ptr(val), csr.shape[0], len(val), int(cplx), 0,
                                              C.byref(err)))
        return err.value

    def reduce(self, what):
        out = np.empty(self.desc.batch, dtype=np.float64)
        self._check(self.lib.nls_reduce(self.h, what, out.ctypes.data_as(_dp)))
        return out

    def slab_export(self):
        buf = C.create_string_buffer(128)
        self._check(self.lib.nls_slab_export(self.h, buf))
        return buf.raw

    def slab_connect(self, blobs, rank, nranks):
        assert len(blobs) == 128 * nranks
        self._check(self.lib.nls_slab_connect(self.h, blobs, rank, nranks))

    def launch_count(self):
        return int(self.lib.nls_launch_count(self.h))

    def tmap_launch_count(self):
        return int(self.lib.nls_tmap_launch_count(self.h))

    def tma_launch_count(self):
        return int(self.lib.nls_tma_launch_count(self.h))
