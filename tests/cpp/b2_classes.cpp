// b2_classes.cpp -- instantiates EVERY class of include/nls/solvers.hpp (the reference's C++ solver-class surface,
// SURVEY.md 8b "B2") through the reference's own constructor signatures and runs the reference's driver loops:
//   nlsolvers/device/drivers/nlse_cubic_driver_2d.cpp:100-116   (gen-2 NLSE: matrix + online snapshots)
//   nlsolvers/device/drivers/kg_driver_dev_2d.cpp:129-157       (gen-2 KGE: device CSR arrays, store_snapshot)
//   device/sg_driver_dev.cpp:87-140                             (CSR upload + solver loop + transfer_snapshots)
// The matrices are assembled by the CPU oracle (test infrastructure; orc_csr = the reference's builders,
// laplacians.hpp:9-218) exactly as the reference drivers assemble them with Eigen, and handed over as
// Eigen::SparseMatrix-like objects / device CSR arrays.  Outputs are .npy files that tests/test_b2_classes.py
// compares with the CLI drivers (bit-equal) and the oracle.  TEST PROGRAM: links the oracle, never shipped.
#include "../../include/nls/solvers.hpp"
#include "../../nonlinear-solvers_b200/host/npy.hpp"
#include "../../oracle/nls_oracle.h"

#include <cuda_runtime.h>

#include <iostream>

using cplx = std::complex<double>;
using nls_host::read_npy;
using nls_host::write_npy;

// the accessors of Eigen::SparseMatrix that the reference's constructors use (compressed storage)
template <class T> struct SparseLike {
  std::vector<int> outer, inner;
  std::vector<T> values;
  int rows() const { return (int)outer.size() - 1; }
  long nonZeros() const { return (long)values.size(); }
  const int *outerIndexPtr() const { return outer.data(); }
  const int *innerIndexPtr() const { return inner.data(); }
  const T *valuePtr() const { return values.data(); }
};

static void ock(int rc, const char *what) {
  if (rc) throw std::runtime_error(std::string(what) + ": " + orc_last_error());
}
static void cuk(cudaError_t e, const char *what) {
  if (e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e));
}

// sign * build_[anisotropic_]laplacian_noflux[_3d](...) as the reference drivers call it
template <class T>
SparseLike<T> assemble(uint32_t nx, uint32_t ny, uint32_t nz, double dx, const double *c, double sign, bool cplx_state) {
  nls_desc d;
  std::memset(&d, 0, sizeof d);
  d.struct_size = sizeof d;
  d.nx = nx; d.ny = ny; d.nz = nz; d.dx = d.dy = d.dz = dx;
  d.equation = cplx_state ? NLS_EQ_NLSE_CUBIC : NLS_EQ_KGE;
  d.integrator = cplx_state ? NLS_INT_SS2 : NLS_INT_GAUTSCHI;
  d.laplacian = c ? NLS_LAP_ANISO : NLS_LAP_UNIFORM;
  d.krylov_m = 10; d.batch = 1; d.op_sign = sign; d.dt = 1e-3;
  orc_handle o;
  ock(orc_create(&d, &o), "orc_create");
  const size_t n = (size_t)nx * ny * nz;
  if (c) ock(orc_set_field(o, NLS_FIELD_C, c, n * 8), "orc_set_field(C)");
  int64_t nnz = 0;
  ock(orc_csr(o, &nnz, nullptr, nullptr, nullptr), "orc_csr");
  std::vector<int64_t> rp(n + 1);
  std::vector<int32_t> col(nnz);
  std::vector<double> val(nnz);
  ock(orc_csr(o, &nnz, rp.data(), col.data(), val.data()), "orc_csr");
  orc_destroy(o);
  SparseLike<T> L;
  L.outer.assign(rp.begin(), rp.end());
  L.inner.assign(col.begin(), col.end());
  L.values.resize(nnz);
  for (int64_t k = 0; k < nnz; ++k) L.values[k] = T(val[k]);
  return L;
}

template <class T> struct DeviceCsr { // what the reference drivers cudaMalloc + cudaMemcpy (kg_driver_dev_2d.cpp:129-139)
  int *row_ptr = nullptr, *col_ind = nullptr;
  T *values = nullptr;
  explicit DeviceCsr(const SparseLike<T> &L) {
    cuk(cudaMalloc(&row_ptr, (L.rows() + 1) * sizeof(int)), "cudaMalloc");
    cuk(cudaMalloc(&col_ind, L.nonZeros() * sizeof(int)), "cudaMalloc");
    cuk(cudaMalloc(&values, L.nonZeros() * sizeof(T)), "cudaMalloc");
    cuk(cudaMemcpy(row_ptr, L.outerIndexPtr(), (L.rows() + 1) * sizeof(int), cudaMemcpyHostToDevice), "cudaMemcpy");
    cuk(cudaMemcpy(col_ind, L.innerIndexPtr(), L.nonZeros() * sizeof(int), cudaMemcpyHostToDevice), "cudaMemcpy");
    cuk(cudaMemcpy(values, L.valuePtr(), L.nonZeros() * sizeof(T), cudaMemcpyHostToDevice), "cudaMemcpy");
  }
  ~DeviceCsr() { cudaFree(row_ptr); cudaFree(col_ind); cudaFree(values); }
};

template <class T> std::vector<T> load(const std::string &dir, const char *name, std::vector<uint64_t> &shape) {
  return read_npy<T>(dir + "/" + name, shape);
}

int main(int argc, char **argv) {
  if (argc != 7) {
    std::cerr << "usage: b2_classes <dir> nx L T nt num_snapshots   (inputs: u0c.npy u0r.npy v0.npy m.npy c.npy [3D: *_3d.npy])\n";
    return 1;
  }
  const std::string dir = argv[1];
  const uint32_t nx = std::stoul(argv[2]), ny = nx;
  const double Lbox = std::stod(argv[3]), T = std::stod(argv[4]);
  const uint32_t nt = std::stoul(argv[5]), num_snapshots = std::stoul(argv[6]);
  const double dx = 2 * Lbox / (nx - 1), dt = T / nt;
  const uint32_t freq = nt / num_snapshots, n = nx * ny;
  const cplx dti(0, dt);
  std::vector<uint64_t> sh;
  const auto u0c = load<cplx>(dir, "u0c.npy", sh);
  const auto u0r = load<double>(dir, "u0r.npy", sh);
  const auto v0 = load<double>(dir, "v0.npy", sh);
  const auto m = load<double>(dir, "m.npy", sh);
  const auto c = load<double>(dir, "c.npy", sh);
  const std::vector<uint64_t> shape = {num_snapshots, ny, nx};

  { // ---- gen-2 NLSE: nlse_cubic_driver_2d.cpp:94-116 (L = build_anisotropic_laplacian_noflux<complex>(...), Krylov 20)
    const auto L = assemble<cplx>(nx, ny, 1, dx, c.data(), 1.0, true);
    std::vector<cplx> u_save((size_t)num_snapshots * n);
    bool is_3d = false;
    device::Parameters params(num_snapshots, freq, 20);
    device::NLSESolverDevice solver(L, u0c.data(), m.data(), is_3d, params);
    solver.store_snapshot_online(u_save.data());
    for (uint32_t i = 1; i < nt; ++i) {
      solver.step(dti, i, u_save.data());
      solver.apply_bc();
    }
    write_npy(dir + "/b2_nlse_gen2.npy", u_save.data(), shape);
    std::cout << "gen-2 NLSESolverDevice: recognised " << solver.grid().nx << "x" << solver.grid().ny
              << (solver.grid().c ? " div(c grad)" : " uniform") << "\n";
  }
  { // ---- gen-2 sEWI: same constructor with sewi = true, step_sewi (nlse_cubic_sewi_driver_2d.cpp, Krylov 25)
    const auto L = assemble<cplx>(nx, ny, 1, dx, c.data(), 1.0, true);
    std::vector<cplx> u_save((size_t)num_snapshots * n);
    device::Parameters params(num_snapshots, freq, 25);
    device::NLSESolverDevice solver(L, u0c.data(), m.data(), false, params, true);
    solver.store_snapshot_online(u_save.data());
    for (uint32_t i = 1; i < nt; ++i) solver.step_sewi(dti, i, u_save.data());
    write_npy(dir + "/b2_nlse_sewi.npy", u_save.data(), shape);
  }
  { // ---- gen-1 NLSE: device/nlse_solver_dev.hpp:53-118 (uniform matrix, no m field, transfer_snapshots)
    const auto L = assemble<cplx>(nx, ny, 1, dx, nullptr, 1.0, true);
    std::vector<cplx> u_save((size_t)num_snapshots * n);
    NLSESolverDevice solver(L, u0c.data(), device::Parameters(num_snapshots, freq, 10));
    for (uint32_t i = 1; i < nt; ++i) solver.step(dti, i);
    solver.transfer_snapshots(u_save.data());
    write_npy(dir + "/b2_nlse_gen1.npy", u_save.data(), shape);
  }
  { // ---- cubic-quintic and saturable classes (nlse_cubic_quintic_dev.hpp:30-105, nlse_saturating_dev.hpp)
    const auto L = assemble<cplx>(nx, ny, 1, dx, nullptr, 1.0, true);
    device::Parameters pq(num_snapshots, freq, 15);
    pq.sigma1 = 1.0; pq.sigma2 = -0.1;
    std::vector<cplx> u_save((size_t)num_snapshots * n);
    device::NLSECubicQuinticSolverDevice cq(L, u0c.data(), m.data(), pq);
    for (uint32_t i = 1; i < nt; ++i) { cq.step(dti, i); cq.apply_bc(); }
    cq.transfer_snapshots(u_save.data());
    write_npy(dir + "/b2_nlse_cq.npy", u_save.data(), shape);
    device::Parameters ps(num_snapshots, freq, 10);
    ps.kappa = 0.3;
    nls::Grid g;
    g.nx = nx; g.ny = ny; g.dx = g.dy = dx;
    device::NLSESaturatingSolverDevice sat(g, u0c.data(), m.data(), ps); // geometry constructor
    for (uint32_t i = 1; i < nt; ++i) { sat.step(dti, i); sat.apply_bc(); }
    sat.transfer_snapshots(u_save.data());
    write_npy(dir + "/b2_nlse_sat.npy", u_save.data(), shape);
  }
  { // ---- gen-2 KGE: kg_driver_dev_2d.cpp:107-157 (L = -build_anisotropic_laplacian_noflux<double>, device CSR arrays)
    const auto L = assemble<double>(nx, ny, 1, dx, c.data(), -1.0, false);
    DeviceCsr<double> d(L);
    std::vector<double> u_save((size_t)num_snapshots * n), v_save((size_t)num_snapshots * n);
    device::Parameters params(num_snapshots, freq, 10);
    bool is_3d = false;
    device::KGESolverDevice solver(d.row_ptr, d.col_ind, d.values, m.data(), nx * ny, (uint32_t)L.nonZeros(), u0r.data(),
                                   v0.data(), dt, is_3d, params);
    for (uint32_t i = 1; i < nt; ++i) {
      solver.step();
      solver.apply_bc();
      if (i % freq == 0) {
        uint32_t snapshot_idx = i / freq;
        if (snapshot_idx < num_snapshots) solver.store_snapshot(snapshot_idx);
      }
    }
    solver.transfer_snapshots(u_save.data(), v_save.data());
    write_npy(dir + "/b2_kge_u.npy", u_save.data(), shape);
    write_npy(dir + "/b2_kge_v.npy", v_save.data(), shape);
    std::cout << "gen-2 KGESolverDevice: recognised op_sign " << solver.grid().op_sign << "\n";
  }
  { // ---- SGE: sg_single_dev.hpp:28-31 constructor (no is_3d), device/sg_driver_dev.cpp:87-140 loop shape
    const auto L = assemble<double>(nx, ny, 1, dx, nullptr, 1.0, false);
    DeviceCsr<double> d(L);
    std::vector<double> u_save((size_t)num_snapshots * n), v_save((size_t)num_snapshots * n);
    SGESolverDevice solver(d.row_ptr, d.col_ind, d.values, m.data(), nx * ny, (uint32_t)L.nonZeros(), u0r.data(), v0.data(), dt,
                           device::Parameters(num_snapshots, freq, 10));
    for (uint32_t i = 1; i < nt; ++i) solver.step(dt, i);
    solver.transfer_snapshots(u_save.data(), 'u');
    solver.transfer_snapshots(v_save.data(), 'v');
    write_npy(dir + "/b2_sge_u.npy", u_save.data(), shape);
    write_npy(dir + "/b2_sge_v.npy", v_save.data(), shape);
    // Stoermer-Verlet through the same class (kg_dev.hpp:93-110 step_sv)
    device::SGESolverDevice sv(d.row_ptr, d.col_ind, d.values, m.data(), nx * ny, (uint32_t)L.nonZeros(), u0r.data(), v0.data(), dt,
                               device::Parameters(num_snapshots, freq, 10), true);
    for (uint32_t i = 1; i < nt; ++i) {
      sv.step_sv();
      if (i % freq == 0 && i / freq < num_snapshots) sv.store_snapshot(i / freq);
    }
    sv.transfer_snapshots(u_save.data(), v_save.data());
    write_npy(dir + "/b2_sge_sv_u.npy", u_save.data(), shape);
  }
  { // ---- the remaining real-equation classes through the geometry constructor
    nls::Grid g;
    g.nx = nx; g.ny = ny; g.dx = g.dy = dx; g.op_sign = -1.0;
    std::vector<double> us(u0r), u_save((size_t)num_snapshots * n);
    for (auto &x : us) x *= 0.2;
    device::Phi4SolverDevice p4(g, m.data(), us.data(), v0.data(), dt, device::Parameters(num_snapshots, freq, 10));
    device::SGEDoubleSolverDevice sd(g, m.data(), us.data(), v0.data(), dt, device::Parameters(num_snapshots, freq, 10));
    device::SGEHyperbolicSolverDevice sh2(g, m.data(), us.data(), v0.data(), dt, device::Parameters(num_snapshots, freq, 10));
    for (uint32_t i = 1; i < nt; ++i) {
      p4.step(); sd.step(); sh2.step();
      if (i % freq == 0 && i / freq < num_snapshots) { p4.store_snapshot(i / freq); sd.store_snapshot(i / freq); sh2.store_snapshot(i / freq); }
    }
    p4.transfer_snapshots(u_save.data());
    write_npy(dir + "/b2_phi4_u.npy", u_save.data(), shape);
    sd.transfer_snapshots(u_save.data());
    write_npy(dir + "/b2_sgdouble_u.npy", u_save.data(), shape);
    sh2.transfer_snapshots(u_save.data());
    write_npy(dir + "/b2_sinh_u.npy", u_save.data(), shape);
  }
  { // ---- matrix-function applicators + SpMV on DEVICE pointers (device/matfunc_complex.hpp:155-156,
    //      matfunc_real.hpp:171, spmv.hpp:65-72), constructed from the reference's matrix arguments
    const auto Lc = assemble<cplx>(nx, ny, 1, dx, c.data(), 1.0, true);
    const auto Lr = assemble<double>(nx, ny, 1, dx, c.data(), -1.0, false);
    DeviceCsr<double> dr(Lr);
    DeviceCsr<cplx> dc(Lc);
    cplx *d_in, *d_out;
    double *d_rin, *d_rout;
    cuk(cudaMalloc(&d_in, n * 16), "cudaMalloc"); cuk(cudaMalloc(&d_out, n * 16), "cudaMalloc");
    cuk(cudaMalloc(&d_rin, n * 8), "cudaMalloc"); cuk(cudaMalloc(&d_rout, n * 8), "cudaMalloc");
    cuk(cudaMemcpy(d_in, u0c.data(), n * 16, cudaMemcpyHostToDevice), "cudaMemcpy");
    cuk(cudaMemcpy(d_rin, u0r.data(), n * 8, cudaMemcpyHostToDevice), "cudaMemcpy");
    std::vector<cplx> hc(n);
    std::vector<double> hr(n);
    const std::vector<uint64_t> s2 = {ny, nx};
    MatrixFunctionApplicatorComplex mc(Lc, n, 10, (uint32_t)Lc.nonZeros());
    mc.apply(d_out, d_in, cplx(0, 1e-2));
    mc.sync();
    cuk(cudaMemcpy(hc.data(), d_out, n * 16, cudaMemcpyDeviceToHost), "cudaMemcpy");
    write_npy(dir + "/b2_matfunc_exp.npy", hc.data(), s2);
    mc.apply(d_out, d_in, cplx(1e-2, 0), "sinc");
    mc.sync();
    cuk(cudaMemcpy(hc.data(), d_out, n * 16, cudaMemcpyDeviceToHost), "cudaMemcpy");
    write_npy(dir + "/b2_matfunc_sinc.npy", hc.data(), s2);
    mc.apply(hc.data(), u0c.data(), cplx(0, 1e-2)); // host pointers: blocking convenience path
    write_npy(dir + "/b2_matfunc_exp_host.npy", hc.data(), s2);
    MatrixFunctionApplicatorReal mr(dr.row_ptr, dr.col_ind, dr.values, n, 10, (uint32_t)Lr.nonZeros());
    const char *fnames[] = {"cos_sqrt", "sinc2_sqrt", "id_sqrt"};
    const FunctionType fts[] = {FunctionType::COS_SQRT, FunctionType::SINC2_SQRT, FunctionType::ID_SQRT};
    for (int k = 0; k < 3; ++k) {
      mr.apply(d_rout, d_rin, 1e-2, fts[k]);
      mr.sync();
      cuk(cudaMemcpy(hr.data(), d_rout, n * 8, cudaMemcpyDeviceToHost), "cudaMemcpy");
      write_npy(dir + "/b2_matfunc_" + fnames[k] + ".npy", hr.data(), s2);
    }
    DeviceSpMV<double> sr(dr.row_ptr, dr.col_ind, dr.values, n, (uint32_t)Lr.nonZeros());
    sr.multiply(d_rin, d_rout);
    sr.sync();
    cuk(cudaMemcpy(hr.data(), d_rout, n * 8, cudaMemcpyDeviceToHost), "cudaMemcpy");
    write_npy(dir + "/b2_spmv_real.npy", hr.data(), s2);
    DeviceSpMV<cplx> sc(dc.row_ptr, dc.col_ind, dc.values, n, (uint32_t)Lc.nonZeros());
    sc.multiply(d_in, d_out);
    sc.sync();
    cuk(cudaMemcpy(hc.data(), d_out, n * 16, cudaMemcpyDeviceToHost), "cudaMemcpy");
    write_npy(dir + "/b2_spmv_cplx.npy", hc.data(), s2);
    cudaFree(d_in); cudaFree(d_out); cudaFree(d_rin); cudaFree(d_rout);
  }
  { // ---- 3D, is_3d = true, div(c grad) with the y-wrap links (even ny pins the recovered c field)
    std::vector<uint64_t> s3;
    const auto u3 = load<double>(dir, "u0r_3d.npy", s3);
    const auto v3 = load<double>(dir, "v0_3d.npy", s3);
    const auto c3 = load<double>(dir, "c_3d.npy", s3);
    const uint32_t n3 = (uint32_t)s3[0];
    const double dx3 = 2 * Lbox / (n3 - 1);
    const auto L = assemble<double>(n3, n3, n3, dx3, c3.data(), -1.0, false);
    DeviceCsr<double> d(L);
    const uint32_t nn = n3 * n3 * n3;
    std::vector<double> ones(nn, 1.0), u_save((size_t)num_snapshots * nn), v_save((size_t)num_snapshots * nn);
    device::KGESolverDevice solver(d.row_ptr, d.col_ind, d.values, ones.data(), nn, (uint32_t)L.nonZeros(), u3.data(), v3.data(), dt,
                                   true, device::Parameters(num_snapshots, freq, 10));
    for (uint32_t i = 1; i < nt; ++i) {
      solver.step();
      if (i % freq == 0 && i / freq < num_snapshots) solver.store_snapshot(i / freq);
    }
    solver.transfer_snapshots(u_save.data(), v_save.data());
    write_npy(dir + "/b2_kge3d_u.npy", u_save.data(), {num_snapshots, n3, n3, n3});
  }
  { // ---- a matrix that is NOT one of the stencils must be rejected (no cuSPARSE / CPU fallback by design)
    auto L = assemble<double>(nx, ny, 1, dx, nullptr, 1.0, false);
    L.values[L.values.size() / 2] *= 1.0 + 1e-6;
    DeviceCsr<double> d(L);
    bool rejected = false;
    try {
      DeviceSpMV<double> bad(d.row_ptr, d.col_ind, d.values, n, (uint32_t)L.nonZeros());
    } catch (const std::runtime_error &e) {
      rejected = true;
      std::cout << "rejected as expected: " << e.what() << "\n";
    }
    if (!rejected) { std::cerr << "a perturbed matrix was accepted\n"; return 2; }
  }
  std::cout << "B2_OK\n";
  return 0;
}
