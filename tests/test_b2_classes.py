"""The B2 class layer (include/nls/solvers.hpp: the reference's C++ solver classes) exercised by a
C++ program (tests/cpp/b2_classes.cpp) that instantiates every class through the reference's constructor
signatures -- Eigen::SparseMatrix-like objects, device CSR arrays, device-pointer apply()/multiply() -- and runs the
reference's driver loops.  Its .npy outputs are compared with the CLI drivers and with the oracle.

Classes built from nls::Grid run the very same handle configuration as the CLI: bit-equal.  Classes built from an
assembled matrix recover spacing / c(x) from the matrix entries (1/sqrt(w), tree propagation of c_p + c_q): the
operator is the same to rounding, so those trajectories agree to ~1e-12 instead of bit for bit."""
import os
import subprocess

import numpy as np
import pytest

from conftest import rel_l2
import ics
import nls_b200
from nls_b200 import Desc, EQ, INT, LAP, FN, FIELD
from test_drivers import oracle_trajectory

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "nonlinear-solvers_b200", "host")
BIN = os.path.join(HOST, "bin")
N, L, T, NT, NS, N3 = 48, 6.0, 0.04, 40, 8, 12


def test_b2_program_builds(oracle_lib):
    """CPU: the header compiles with every class instantiated and links against the C ABI"""
    nls_b200.build_library()
    subprocess.run(["make", "-C", HOST, "-s", "bin/b2_classes"], check=True)
    assert os.path.exists(os.path.join(BIN, "b2_classes"))


@pytest.fixture(scope="module")
def b2(tmp_path_factory, oracle_lib):
    nls_b200.build_library()
    subprocess.run(["make", "-C", HOST, "-s"], check=True)
    subprocess.run(["make", "-C", HOST, "-s", "bin/b2_classes"], check=True)
    d = tmp_path_factory.mktemp("b2")
    rng = np.random.default_rng(3)
    x = np.linspace(-L, L, N)
    X, Y = np.meshgrid(x, x)
    inp = {
        "u0c": ics.gaussian_bumps_2d(N, L),
        "u0r": 0.3 * ics.kink_2d(N, L),
        "m": 1.0 + 0.2 * np.sin(X) * np.cos(0.5 * Y),
        "c": 1.0 + 0.3 * np.exp(-(X ** 2 + Y ** 2) / 8.0) + 0.05 * rng.standard_normal((N, N)),
    }
    inp["v0"] = 0.05 * np.cos(inp["u0r"])
    x3 = np.linspace(-L, L, N3)
    Z3, Y3, X3 = np.meshgrid(x3, x3, x3, indexing="ij")
    inp["u0r_3d"] = 0.3 * np.abs(ics.gaussian_bumps_3d(N3, L))
    inp["v0_3d"] = 0.05 * np.cos(inp["u0r_3d"])
    inp["c_3d"] = 1.0 + 0.3 * np.exp(-(X3 ** 2 + Y3 ** 2 + Z3 ** 2) / 8.0) + 0.05 * rng.standard_normal((N3, N3, N3))
    for k, v in inp.items():
        np.save(d / f"{k}.npy", v)
    np.save(d / "ones.npy", np.ones((N, N)))
    np.save(d / "u0r_small.npy", 0.2 * inp["u0r"])
    r = subprocess.run([os.path.join(BIN, "b2_classes"), str(d), str(N), str(L), str(T), str(NT), str(NS)],
                       capture_output=True, text=True)
    print(r.stdout[-3000:], r.stderr[-3000:])
    assert r.returncode == 0 and "B2_OK" in r.stdout and "rejected as expected" in r.stdout
    return d, inp


def cli(name, *args):
    subprocess.run([os.path.join(BIN, name)] + [str(a) for a in args], check=True, capture_output=True, text=True)


def out(d, name):
    return np.load(d / f"b2_{name}.npy")


@pytest.mark.gpu
def test_gen2_nlse_matrix_constructor_runs_the_reference_loop(b2, oracle_lib):
    d, inp = b2
    got = out(d, "nlse_gen2")
    cli("nlse_2d_dev", N, N, L, L, d / "u0c.npy", d / "cli_nlse2.npy", T, NT, NS, d / "m.npy", d / "c.npy")
    dd = Desc.make(N, N, L=L, equation=EQ.NLSE_CUBIC, laplacian=LAP.ANISO, krylov_m=20)
    ref, _ = oracle_trajectory(oracle_lib, dd, inp["u0c"], None, inp["m"].ravel(), inp["c"].ravel(), T, NT, NS)
    e_cli, e_orc = rel_l2(got, np.load(d / "cli_nlse2.npy")), rel_l2(got, ref)
    print(f"gen-2 NLSE class vs CLI {e_cli:.3e}, vs oracle {e_orc:.3e}")
    assert np.array_equal(got[0], inp["u0c"]) and e_cli < 1e-11 and e_orc < 1e-11


@pytest.mark.gpu
def test_gen2_sewi_and_gen1_nlse(b2, oracle_lib):
    d, inp = b2
    cli("nlse_sewi_2d_dev", N, N, L, L, d / "u0c.npy", d / "cli_sewi.npy", T, NT, NS, d / "m.npy", d / "c.npy")
    assert rel_l2(out(d, "nlse_sewi"), np.load(d / "cli_sewi.npy")) < 1e-11
    cli("nlse_cubic", N, N, L, L, d / "u0c.npy", d / "cli_gen1.npy", T, NT, NS, d / "ones.npy")
    got = out(d, "nlse_gen1")
    assert rel_l2(got, np.load(d / "cli_gen1.npy")) < 1e-12
    dd = Desc.make(N, N, L=L, equation=EQ.NLSE_CUBIC, krylov_m=10)
    ref, _ = oracle_trajectory(oracle_lib, dd, inp["u0c"], None, np.ones(N * N), None, T, NT, NS)
    assert rel_l2(got, ref) < 1e-11


@pytest.mark.gpu
def test_cq_and_saturating_classes(b2):
    d, inp = b2
    cli("nlse_cubic_quintic_dev", N, N, L, L, 1.0, -0.1, d / "u0c.npy", d / "cli_cq.npy", T, NT, NS, d / "m.npy")
    assert rel_l2(out(d, "nlse_cq"), np.load(d / "cli_cq.npy")) < 1e-12
    cli("nlse_saturating", N, N, L, L, 0.3, d / "u0c.npy", d / "cli_sat.npy", T, NT, NS, d / "m.npy")
    assert np.array_equal(out(d, "nlse_sat"), np.load(d / "cli_sat.npy")), "geometry constructor: same handle as the CLI"


@pytest.mark.gpu
def test_real_solver_classes_from_device_csr(b2, oracle_lib):
    d, inp = b2
    cli("kg_gautschi_2d_dev", N, N, L, L, d / "u0r.npy", d / "v0.npy", d / "cli_kgu.npy", d / "cli_kgv.npy", T, NT, NS,
        d / "m.npy", d / "c.npy")
    e_u, e_v = rel_l2(out(d, "kge_u"), np.load(d / "cli_kgu.npy")), rel_l2(out(d, "kge_v"), np.load(d / "cli_kgv.npy"))
    dd = Desc.make(N, N, L=L, equation=EQ.KGE, integrator=INT.GAUTSCHI, laplacian=LAP.ANISO, krylov_m=10, dt=T / NT, op_sign=-1.0)
    ru, rv = oracle_trajectory(oracle_lib, dd, inp["u0r"], inp["v0"], inp["m"].ravel(), inp["c"].ravel(), T, NT, NS)
    print(f"KGE class (device CSR, -L, div(c grad)) vs CLI u {e_u:.3e} v {e_v:.3e}; vs oracle {rel_l2(out(d, 'kge_u'), ru):.3e}")
    assert e_u < 1e-11 and e_v < 1e-8 and rel_l2(out(d, "kge_u"), ru) < 1e-11
    cli("sg_single", N, N, L, L, d / "u0r.npy", d / "v0.npy", d / "cli_sgu.npy", d / "cli_sgv.npy", T, NT, NS, d / "m.npy")
    assert rel_l2(out(d, "sge_u"), np.load(d / "cli_sgu.npy")) < 1e-12
    assert rel_l2(out(d, "sge_v"), np.load(d / "cli_sgv.npy")) < 1e-8
    cli("sg_single_sv", N, N, L, L, d / "u0r.npy", d / "v0.npy", d / "cli_svu.npy", d / "cli_svv.npy", T, NT, NS, d / "m.npy")
    assert rel_l2(out(d, "sge_sv_u"), np.load(d / "cli_svu.npy")) < 1e-12


@pytest.mark.gpu
def test_geometry_constructed_real_classes_are_bit_equal_to_the_cli(b2):
    d, inp = b2
    for name, exe in (("phi4_u", "phi4"), ("sgdouble_u", "sg_double"), ("sinh_u", "sg_hyperbolic")):
        cli(exe, N, N, L, L, d / "u0r_small.npy", d / "v0.npy", d / f"cli_{name}.npy", d / "cli_tmp_v.npy", T, NT, NS, d / "m.npy")
        assert np.array_equal(out(d, name), np.load(d / f"cli_{name}.npy")), name


@pytest.mark.gpu
def test_applicators_and_spmv_on_device_pointers(b2, oracle_lib):
    d, inp = b2
    dc = Desc.make(N, N, L=L, equation=EQ.NLSE_CUBIC, laplacian=LAP.ANISO, krylov_m=10)
    oc = oracle_lib.Oracle(dc)
    oc.set(FIELD.C, inp["c"].ravel())
    assert rel_l2(out(d, "matfunc_exp"), oc.matfunc(FN.EXP, 1e-2j, inp["u0c"].ravel())) < 1e-11
    assert rel_l2(out(d, "matfunc_exp_host"), out(d, "matfunc_exp")) == 0.0
    assert rel_l2(out(d, "matfunc_sinc"), oc.matfunc(FN.SINC, 1e-2, inp["u0c"].ravel())) < 1e-11
    assert rel_l2(out(d, "spmv_cplx"), oc.stencil_apply(inp["u0c"].ravel())) < 1e-12
    dr = Desc.make(N, N, L=L, equation=EQ.KGE, integrator=INT.GAUTSCHI, laplacian=LAP.ANISO, krylov_m=10, op_sign=-1.0)
    orr = oracle_lib.Oracle(dr)
    orr.set(FIELD.C, inp["c"].ravel())
    for name, fn in (("cos_sqrt", FN.COS_SQRT), ("sinc2_sqrt", FN.SINC2_SQRT), ("id_sqrt", FN.ID_SQRT)):
        assert rel_l2(out(d, f"matfunc_{name}"), orr.matfunc(fn, 1e-2, inp["u0r"].ravel())) < 1e-11, name
    assert rel_l2(out(d, "spmv_real"), orr.stencil_apply(inp["u0r"].ravel())) < 1e-12


@pytest.mark.gpu
def test_kge_3d_class_recovers_c_through_the_wrap_links(b2, oracle_lib):
    d, inp = b2
    dd = Desc.make(N3, N3, N3, L=L, equation=EQ.KGE, integrator=INT.GAUTSCHI, laplacian=LAP.ANISO, krylov_m=10, dt=T / NT,
                   op_sign=-1.0)
    ru, _ = oracle_trajectory(oracle_lib, dd, inp["u0r_3d"], inp["v0_3d"], np.ones(N3 ** 3), inp["c_3d"].ravel(), T, NT, NS)
    err = rel_l2(out(d, "kge3d_u"), ru)
    print(f"3D KGE class (is_3d, device CSR) vs oracle {err:.3e}")
    assert err < 1e-11


@pytest.mark.gpu
def test_csr_residual_abi(oracle_lib):
    """nls_csr_residual: ~1e-16 for the handle's own operator, O(1) for a different one"""
    n = 40
    for lap in (LAP.UNIFORM, LAP.ANISO):
        d = Desc.make(n, n, L=5.0, equation=EQ.NLSE_CUBIC, laplacian=lap)
        c = 1.0 + 0.2 * np.random.default_rng(4).standard_normal(d.n)
        e, o = nls_b200.Engine(d), oracle_lib.Oracle(d)
        if lap == LAP.ANISO:
            e.set(FIELD.C, c); o.set(FIELD.C, c)
        A = o.csr()
        assert e.csr_residual(A) < 1e-14
        assert e.csr_residual(A.astype(np.complex128)) < 1e-14
        B = A.copy()
        B.data[B.nnz // 3] *= 1.001
        assert e.csr_residual(B) > 1e-6
