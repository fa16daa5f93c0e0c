"""CPU tests of the drop-in boundary: the C-ABI library builds for sm_100a, loads, exports every
symbol include/nls.h declares, and fails loudly (no CPU fallback) when no GPU is present."""
import ctypes as C
import os
import re

import pytest

import nls_b200
from nls_b200 import Desc, EQ
from nls_b200 import binding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    nls_b200.build_library()
    return nls_b200.load_library()


def test_header_symbols_all_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "nls.h")).read()
    declared = set(re.findall(r"\b(nls_[a-z_0-9]+)\s*\(", hdr)) - {"nls_allreduce_fn"}
    assert declared, "no declarations found"
    raw = C.CDLL(nls_b200.library_path())
    for name in sorted(declared):
        assert hasattr(raw, name), f"{name} declared in nls.h but not exported"
    assert declared == set(binding.EXPORTS), declared ^ set(binding.EXPORTS)


def test_desc_layout_matches_header(lib):
    # nls_create rejects a descriptor whose struct_size differs from sizeof(nls_desc)
    d = Desc.make(8, 8, L=1.0)
    d.struct_size += 4
    h = C.c_void_p()
    assert lib.nls_create(C.byref(d), C.byref(h)) == -1
    assert b"descriptor" in lib.nls_last_error()


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    d = Desc.make(16, 16, L=1.0, equation=EQ.NLSE_CUBIC)
    with pytest.raises(nls_b200.NlsError):
        nls_b200.Engine(d)


def test_sass_is_sm100(lib):
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", nls_b200.library_path()], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_noalloc_entry_point_validates_before_touching_cuda(lib):
    """nls_reduce_mass_noalloc rejects sizes beyond its static scratch with NLS_ERR_ARG and a message -- checked before
    any CUDA call, so this runs without a GPU (the compute path: tests/test_zz_no_alloc.py, -m gpu)"""
    buf = (C.c_double * 8)()
    p = C.cast(buf, C.c_void_p)
    assert lib.nls_reduce_mass_noalloc(0, p, 4, 257, 0, 0, 1.0, p, None) == -1
    assert b"NLS_NOALLOC_MAX_BATCH" in lib.nls_last_error()
    assert lib.nls_reduce_mass_noalloc(0, p, (1 << 31) + 1, 1, 0, 1, 1.0, p, None) == -1
    assert b"NLS_NOALLOC_MAX_SITES" in lib.nls_last_error()
    assert lib.nls_reduce_mass_noalloc(0, p, 16, 2, 8, 0, 1.0, p, None) == -1  # traj_stride < n_sites
    assert lib.nls_reduce_mass_noalloc(0, None, 4, 1, 0, 0, 1.0, p, None) == -1
    hdr = open(os.path.join(ROOT, "include", "nls.h")).read()
    assert "NLS_NOALLOC_MAX_BATCH 256u" in hdr and "no counterpart in the reference" in hdr.lower()
