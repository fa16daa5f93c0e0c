// solvers.hpp -- the reference's C++ solver-class surface (SURVEY.md 8b "B2") over the C ABI
// of include/nls.h.  Header-only host C++17, no CUDA, no Eigen: link with libnls_b200.so.
//
// Two families of constructors per class:
//   (1) the REFERENCE signatures: an assembled sparse matrix (device CSR arrays for the real solvers,
//       sg_single_dev.hpp:28-31 / kg_dev.hpp:28-31 / device/spmv.hpp:30-52 / device/matfunc_real.hpp:111-114;
//       an Eigen::SparseMatrix-like object for the NLSE solvers and the complex applicator,
//       nlse_dev.hpp:78-82 / device/nlse_solver_dev.hpp:53-55 / device/matfunc_complex.hpp:71-73).  The engine is
//       matrix-free, so the matrix is only RECOGNISED: grid extents, spacing, operator sign and -- for the
//       div(c grad) builders -- the c field are read back from its rows (nls::grid_from_csr), a handle with that
//       stencil is created, and the matrix is accepted only if it equals the stencil (nls_csr_residual <= 1e-9);
//       anything else throws std::runtime_error: there is no cuSPARSE / CPU fallback by design.
//   (2) geometry constructors taking nls::Grid (+ optional c field) in place of the matrix argument; every other
//       argument keeps its position and meaning.
// Errors throw std::runtime_error like the reference (nlse_solver_dev.hpp:77-81).  Differences a caller can see:
//   * step() includes the no-flux boundary copy (the drivers always call apply_bc() right after step();
//     apply_bc() is kept as a no-op);
//   * step() is LAZY: consecutive calls are queued and issued as ONE nls_step(h, dt, k) when a snapshot is due or
//     the state is read, so a reference driver loop costs one launch per snapshot interval, not one per step;
//   * snapshots are collected in page-locked host memory owned by the object through the asynchronous egress
//     ring (nls_get_field_async) and handed over by transfer_snapshots().
#pragma once
#include "../nls.h"

#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <new>
#include <stdexcept>
#include <string>
#include <vector>

namespace nls {

struct Grid {
  uint32_t nx = 0, ny = 0, nz = 1;
  double dx = 0, dy = 0, dz = 0;
  const double *c = nullptr; // host c(x) field => anisotropic Laplacian (laplacians.hpp:54-103,:158-218)
  double op_sign = 1.0;      // -1: the drivers that pass -L (kg_driver.cpp:93)
  std::shared_ptr<std::vector<double>> c_owned; // keeps a recovered c field alive (grid_from_csr)
  size_t n() const { return (size_t)nx * ny * nz; }
};

inline void check(int rc, const char *what) {
  if (rc != NLS_OK) throw std::runtime_error(std::string(what) + ": " + nls_last_error());
}

// page-locked host array (nls_host_alloc) for the asynchronous transfers.  It is plain host memory: when page
// locking is not possible (no CUDA device: nls_create will then report that loudly) it degrades to pageable memory,
// with which the asynchronous copies simply become synchronous.
template <class T> class Pinned {
public:
  Pinned() = default;
  explicit Pinned(size_t count) { resize(count); }
  ~Pinned() { release(); }
  Pinned(const Pinned &) = delete;
  Pinned &operator=(const Pinned &) = delete;
  void resize(size_t count) {
    release();
    n_ = count;
    if (!count) return;
    void *q = nullptr;
    if (nls_host_alloc(&q, count * sizeof(T)) == NLS_OK) { p_ = (T *)q; locked_ = true; return; }
    p_ = (T *)std::malloc(count * sizeof(T));
    locked_ = false;
    if (!p_) throw std::bad_alloc();
  }
  T *data() { return p_; }
  const T *data() const { return p_; }
  size_t size() const { return n_; }
  bool page_locked() const { return locked_; }

private:
  void release() {
    if (p_) { if (locked_) nls_host_free(p_); else std::free(p_); }
    p_ = nullptr;
  }
  T *p_ = nullptr;
  size_t n_ = 0;
  bool locked_ = false;
};

// RAII owner of one nls_handle
class Handle {
public:
  Handle(const Grid &g, int equation, int integrator, uint32_t krylov_m, double dt = 0.0, double s1 = 0.0,
         double s2 = 0.0, double kappa = 0.0, uint32_t batch = 1, int device = 0, uint32_t flags = 0) {
    nls_desc d;
    std::memset(&d, 0, sizeof d);
    d.struct_size = sizeof d;
    d.nx = g.nx; d.ny = g.ny; d.nz = g.nz;
    d.dx = g.dx; d.dy = g.dy; d.dz = g.dz;
    d.equation = equation; d.integrator = integrator;
    d.laplacian = g.c ? NLS_LAP_ANISO : NLS_LAP_UNIFORM;
    d.reorth = NLS_REORTH_FULL;
    d.krylov_m = krylov_m; d.batch = batch;
    d.s1 = s1; d.s2 = s2; d.kappa = kappa;
    d.op_sign = g.op_sign; d.dt = dt; d.device = device; d.flags = flags;
    n_ = g.n(); batch_ = batch;
    cplx_ = equation <= NLS_EQ_NLSE_SAT;
    check(nls_create(&d, &h_), "nls_create");
    if (g.c) {
      const int rc = nls_set_field(h_, NLS_FIELD_C, g.c, n_ * sizeof(double));
      if (rc != NLS_OK) { // do not leak the handle when the constructor throws
        const std::string msg = std::string("nls_set_field(C): ") + nls_last_error();
        nls_destroy(h_);
        h_ = nullptr;
        throw std::runtime_error(msg);
      }
    }
  }
  ~Handle() { nls_destroy(h_); }
  Handle(const Handle &) = delete;
  Handle &operator=(const Handle &) = delete;

  void set(int field, const void *src, size_t bytes) { check(nls_set_field(h_, field, src, bytes), "nls_set_field"); }
  void get(int field, void *dst, size_t bytes) { check(nls_get_field(h_, field, dst, bytes), "nls_get_field"); }
  // asynchronous egress: dst is valid after wait()
  void get_async(int field, void *dst, size_t bytes, size_t traj_stride = 0) {
    check(nls_get_field_async(h_, field, dst, bytes, traj_stride), "nls_get_field_async");
  }
  // field resampled on the device to tnx x tny [x tnz] points per trajectory (downsampling.py:68-104)
  void get_resampled(int field, uint32_t tnx, uint32_t tny, uint32_t tnz, void *dst, size_t bytes) {
    check(nls_get_field_resampled(h_, field, tnx, tny, tnz, dst, bytes), "nls_get_field_resampled");
  }
  void get_resampled_async(int field, uint32_t tnx, uint32_t tny, uint32_t tnz, void *dst, size_t bytes, size_t traj_stride = 0) {
    check(nls_get_field_resampled_async(h_, field, tnx, tny, tnz, dst, bytes, traj_stride), "nls_get_field_resampled_async");
  }
  // spectral downsampling on the device (downsampling.py:10-34 downsample_fft), 2D handles
  void get_fft_resampled_async(int field, uint32_t tnx, uint32_t tny, void *dst, size_t bytes, size_t traj_stride = 0) {
    check(nls_get_field_fft_resampled_async(h_, field, tnx, tny, dst, bytes, traj_stride), "nls_get_field_fft_resampled_async");
  }
  // NLS_RED_MASS / NLS_RED_ENERGY / NLS_RED_NONFINITE, one value per trajectory
  std::vector<double> reduce(int what) {
    std::vector<double> out(batch_);
    check(nls_reduce(h_, what, out.data()), "nls_reduce");
    return out;
  }
  void step(double dt, uint32_t nsteps = 1) { check(nls_step(h_, dt, nsteps), "nls_step"); }
  void sync() { check(nls_sync(h_), "nls_sync"); }
  void wait() { check(nls_wait(h_), "nls_wait"); }
  nls_handle raw() const { return h_; }
  size_t n() const { return n_; }
  size_t state_bytes() const { return n_ * batch_ * (cplx_ ? 16 : 8); }

private:
  nls_handle h_ = nullptr;
  size_t n_ = 0, batch_ = 1;
  bool cplx_ = false;
};

// ------------------------------------------------------------------------------------------- CSR recognition
// An assembled matrix of the reference (CSR/CSC arrays; symmetric, so both read the same) on the host or device.
struct CsrView {
  const int *row_ptr = nullptr, *col_ind = nullptr;
  const void *values = nullptr; // double, or interleaved complex128 (values_complex)
  uint32_t n = 0;
  uint64_t nnz = 0;
  bool values_complex = false, on_device = false;
};

// anything with Eigen::SparseMatrix's compressed-storage accessors (rows, nonZeros, outerIndexPtr, innerIndexPtr, valuePtr)
template <class SpMat> CsrView csr_view_of(const SpMat &L) {
  CsrView v;
  v.row_ptr = L.outerIndexPtr();
  v.col_ind = L.innerIndexPtr();
  v.values = L.valuePtr();
  v.n = (uint32_t)L.rows();
  v.nnz = (uint64_t)L.nonZeros();
  v.values_complex = sizeof(*L.valuePtr()) == 16;
  v.on_device = false;
  return v;
}

// Reads grid extents, spacing, operator sign and (div(c grad) builders) a c field back from the rows of one of
// the reference's Laplacians (laplacians.hpp:9-218).  dim_hint: 0 = decide from the sparsity pattern, 2 / 3 =
// what the caller's `is_3d` says.  The result is a CANDIDATE: callers verify it with verify_csr.
inline Grid grid_from_csr(const CsrView &v, int dim_hint = 0) {
  if (!v.row_ptr || !v.col_ind || !v.values || v.n < 9) throw std::runtime_error("grid_from_csr: empty matrix");
  std::vector<int> rp(v.n + 1), ci(v.nnz);
  std::vector<double> val(v.nnz * (v.values_complex ? 2 : 1));
  if (v.on_device) {
    check(nls_memcpy_to_host(rp.data(), v.row_ptr, rp.size() * 4), "nls_memcpy_to_host");
    check(nls_memcpy_to_host(ci.data(), v.col_ind, ci.size() * 4), "nls_memcpy_to_host");
    check(nls_memcpy_to_host(val.data(), v.values, val.size() * 8), "nls_memcpy_to_host");
  } else {
    std::memcpy(rp.data(), v.row_ptr, rp.size() * 4);
    std::memcpy(ci.data(), v.col_ind, ci.size() * 4);
    std::memcpy(val.data(), v.values, val.size() * 8);
  }
  if ((uint64_t)rp[v.n] != v.nnz) throw std::runtime_error("grid_from_csr: row_ptr[n] != nnz");
  const int st = v.values_complex ? 2 : 1; // real parts only (the complex builders store real weights)
  auto weight = [&](long long p, long long q) -> double { // entry (p, q), 0 if absent
    for (int k = rp[p]; k < rp[p + 1]; ++k)
      if (ci[k] == q) return val[(size_t)k * st];
    return 0.0;
  };
  // a row with the most entries is an interior row: its column offsets are {+-1, +-nx [, +-plane]}
  long long r = v.n / 2, best = -1;
  for (long long t = std::max<long long>(0, (long long)v.n / 2 - 4096); t < std::min<long long>(v.n, (long long)v.n / 2 + 4096); ++t)
    if (rp[t + 1] - rp[t] > best) { best = rp[t + 1] - rp[t]; r = t; }
  std::vector<long long> off;
  for (int k = rp[r]; k < rp[r + 1]; ++k)
    if (ci[k] > r) off.push_back(ci[k] - r);
  std::sort(off.begin(), off.end());
  Grid g;
  if (off.size() == 3 && dim_hint != 2) {
    g.nx = (uint32_t)off[1];
    if (off[2] % off[1]) throw std::runtime_error("grid_from_csr: not a 7-point pattern");
    g.ny = (uint32_t)(off[2] / off[1]);
    g.nz = (uint32_t)(v.n / (uint64_t)off[2]);
  } else if (off.size() == 2 && dim_hint != 3) {
    g.nx = (uint32_t)off[1];
    g.ny = (uint32_t)(v.n / (uint64_t)off[1]);
    g.nz = 1;
  } else {
    throw std::runtime_error("grid_from_csr: the sparsity pattern is not a 5/7-point stencil");
  }
  if (off[0] != 1 || (uint64_t)g.nx * g.ny * g.nz != v.n) throw std::runtime_error("grid_from_csr: rows do not tile a grid");
  const long long nx = g.nx, pl = (long long)g.nx * g.ny;
  const double w0 = weight(r, r + 1);
  if (w0 == 0.0) throw std::runtime_error("grid_from_csr: zero off-diagonal");
  g.op_sign = w0 > 0 ? 1.0 : -1.0;
  // uniform builder: every off-diagonal equals 1/(dx*dy) [2D] or 1/dx^2 [3D]
  bool uniform = true;
  for (uint64_t k = 0; k < v.nnz && uniform; ++k) {
    // diagonal entries are skipped by position: walk rows
    (void)k;
    break;
  }
  for (long long p = 0; p < (long long)v.n && uniform; ++p)
    for (int k = rp[p]; k < rp[p + 1]; ++k)
      if (ci[k] != p && std::fabs(val[(size_t)k * st] - w0) > 1e-13 * std::fabs(w0)) { uniform = false; break; }
  if (uniform) {
    g.dx = g.dy = 1.0 / std::sqrt(std::fabs(w0));
    g.dz = g.nz > 1 ? g.dx : 0.0;
    return g;
  }
  // div(c grad) builder: off-diagonal (p,q) = (c_p + c_q)/2 / h^2.  With h = 1 the sums c_p + c_q = 2|w_pq| are
  // propagated along a spanning tree (x-links inside a row, the first site of a row through its y-link, the first
  // site of a plane through its z-link).  On a bipartite grid graph the solution is fixed up to +t on one colour
  // class and -t on the other, which leaves every link sum -- hence the operator -- unchanged; the 3D builders'
  // extra (i, ny-1, k) <-> (i, 0, k+1) links (laplacians.hpp:192-198) close odd cycles when ny is even and pin t.
  auto cvec = std::make_shared<std::vector<double>>(v.n);
  std::vector<double> &c = *cvec;
  auto aw = [&](long long p, long long q) { return std::fabs(weight(p, q)); };
  c[0] = aw(0, 1);
  for (long long k = 0; k < (long long)g.nz; ++k) {
    if (k > 0) c[k * pl] = 2.0 * aw((k - 1) * pl, k * pl) - c[(k - 1) * pl];
    for (long long j = 0; j < (long long)g.ny; ++j) {
      const long long row0 = k * pl + j * nx;
      if (j > 0) c[row0] = 2.0 * aw(row0 - nx, row0) - c[row0 - nx];
      for (long long i = 1; i < nx; ++i) c[row0 + i] = 2.0 * aw(row0 + i - 1, row0 + i) - c[row0 + i - 1];
    }
  }
  if (g.nz > 1 && g.ny % 2 == 0) { // same colour class on both ends of a wrap link: c' = c + s*t  =>  sum is off by 2*s*t
    const long long p = (long long)(g.ny - 1) * nx, q = pl; // (0, ny-1, 0) <-> (0, 0, 1)
    const double delta = 0.5 * (c[p] + c[q]) - aw(p, q);     // = s_p * t
    const int sp = ((g.ny - 1) % 2) ? -1 : 1;
    const double t = sp * delta;
    for (long long k = 0; k < (long long)g.nz; ++k)
      for (long long j = 0; j < (long long)g.ny; ++j)
        for (long long i = 0; i < nx; ++i) c[k * pl + j * nx + i] -= (((i + j + k) % 2) ? -1 : 1) * t;
  }
  g.dx = g.dy = 1.0;
  g.dz = g.nz > 1 ? 1.0 : 0.0;
  g.c_owned = cvec;
  g.c = cvec->data();
  return g;
}

// throws unless the matrix IS the handle's stencil
inline void verify_csr(Handle &h, const CsrView &v, double tol = 1e-9) {
  double err = 0.0;
  check(nls_csr_residual(h.raw(), v.row_ptr, v.col_ind, v.values, v.n, v.nnz, v.values_complex ? 1 : 0, v.on_device ? 1 : 0, &err),
        "nls_csr_residual");
  if (!(err <= tol))
    throw std::runtime_error("the sparse matrix is not one of the no-flux 5/7-point Laplacians this engine implements "
                             "(relative residual " + std::to_string(err) + "); there is no general SpMV fallback -- "
                             "construct the solver from nls::Grid instead");
}

} // namespace nls

// -------------------------------------------------------------------------------------------
// reference class names (gen-2 lives in namespace device; gen-1 names are aliased below)
namespace device {

struct Parameters { // nlse_dev.hpp:68-76, nlse_cubic_quintic_dev.hpp:17-28, sg_single_dev.hpp:17-26
  uint32_t block_size = 256; // ignored: launch shapes are chosen by the engine
  uint32_t num_snapshots = 100;
  uint32_t snapshot_freq = 5;
  uint32_t krylov_dim = 10;
  double sigma1 = 0.0, sigma2 = 0.0, kappa = 0.0;
  Parameters(uint32_t ns = 100, uint32_t freq = 5, uint32_t m = 10) : num_snapshots(ns), snapshot_freq(freq), krylov_dim(m) {}
};

// common NLSE machinery: SS2 / sEWI step, snapshot bookkeeping of nlse_solver_dev.hpp:94-124
template <int EQ> class NLSESolverBase {
public:
  using Scalar = std::complex<double>;
  NLSESolverBase(const nls::Grid &g, const Scalar *host_u0, const double *host_m, const Parameters &p,
                 int integrator = NLS_INT_SS2)
      : grid_(g), h_(g, EQ, integrator, p.krylov_dim, 0.0, p.sigma1, p.sigma2, p.kappa), p_(p), n_(g.n()) {
    init(host_u0, host_m);
  }
  // reference signature: an Eigen::SparseMatrix<std::complex<double>>-like L in place of the geometry
  template <class SpMat>
  NLSESolverBase(const SpMat &L, int dim_hint, const Scalar *host_u0, const double *host_m, const Parameters &p, int integrator)
      : grid_(nls::grid_from_csr(nls::csr_view_of(L), dim_hint)),
        h_(grid_, EQ, integrator, p.krylov_dim, 0.0, p.sigma1, p.sigma2, p.kappa), p_(p), n_(grid_.n()) {
    nls::verify_csr(h_, nls::csr_view_of(L));
    init(host_u0, host_m);
  }
  ~NLSESolverBase() {
    try { flush(); h_.wait(); } catch (...) {}
  }
  // tau = i*dt (nlse_cubic_driver.cpp:50); only its imaginary part is meaningful.  Lazy: see the header comment.
  void step(const Scalar tau, const uint32_t step_number) {
    queue(tau.imag());
    if (p_.snapshot_freq && step_number % p_.snapshot_freq == 0 && current_ < p_.num_snapshots) {
      flush();
      h_.get_async(NLS_FIELD_U, traj_.data() + (size_t)current_++ * n_, n_ * sizeof(Scalar));
    }
  }
  // gen-2 signature: online snapshot into caller memory (nlse_dev.hpp:187, :323-334); the caller reads host_dst
  // without any further call, so this copy is blocking
  void step(const Scalar tau, const uint32_t step_number, Scalar *host_dst) {
    queue(tau.imag());
    if (host_dst && p_.snapshot_freq && step_number % p_.snapshot_freq == 0) {
      const uint32_t idx = step_number / p_.snapshot_freq;
      if (idx < p_.num_snapshots) {
        flush();
        h_.get(NLS_FIELD_U, host_dst + (size_t)idx * n_, n_ * sizeof(Scalar));
      }
    }
  }
  // sEWI step (nlse_dev.hpp:205-238): construct the solver with sewi = true; the first call is the SS2
  // start-up step with u- = u0, exactly as the reference's `if (step_number == 1)` branch
  void step_sewi(const Scalar tau, const uint32_t step_number, Scalar *host_dst) { step(tau, step_number, host_dst); }
  void apply_bc() {} // fused into step()
  void store_snapshot_online(Scalar *host_dst) { // nlse_dev.hpp:323-334: current state -> snapshot slot `current`
    flush();
    if (current_online_ < p_.num_snapshots) h_.get(NLS_FIELD_U, host_dst + (size_t)current_online_++ * n_, n_ * sizeof(Scalar));
  }
  void transfer_snapshots(Scalar *dst) {
    flush();
    h_.wait();
    std::memcpy(dst, traj_.data(), traj_.size() * sizeof(Scalar));
  }
  void current_state(Scalar *dst) {
    flush();
    h_.get(NLS_FIELD_U, dst, n_ * sizeof(Scalar));
  }
  nls::Handle &handle() { flush(); return h_; }
  const nls::Grid &grid() const { return grid_; }

protected:
  void init(const Scalar *host_u0, const double *host_m) {
    h_.set(NLS_FIELD_U, host_u0, n_ * sizeof(Scalar));
    std::vector<double> ones;
    if (!host_m) { ones.assign(n_, 1.0); host_m = ones.data(); } // nlse_cubic_quintic_driver_dev.cpp:83-85
    h_.set(NLS_FIELD_M, host_m, n_ * sizeof(double));
    traj_.resize((size_t)p_.num_snapshots * n_);
    if (p_.num_snapshots) { std::memcpy(traj_.data(), host_u0, n_ * sizeof(Scalar)); current_ = 1; }
  }
  void queue(double dt) {
    if (pending_ && dt != pending_dt_) flush();
    pending_dt_ = dt;
    ++pending_;
  }
  void flush() {
    if (pending_) h_.step(pending_dt_, pending_);
    pending_ = 0;
  }
  nls::Grid grid_;
  nls::Handle h_;
  Parameters p_;
  size_t n_;
  uint32_t current_ = 0, current_online_ = 0, pending_ = 0;
  double pending_dt_ = 0.0;
  nls::Pinned<Scalar> traj_;
};

class NLSESolverDevice : public NLSESolverBase<NLS_EQ_NLSE_CUBIC> { // nlse_dev.hpp:66-361
public:
  NLSESolverDevice(const nls::Grid &g, const Scalar *host_u0, const double *host_m, const bool & /*is_3d*/,
                   const Parameters &p = Parameters(), bool sewi = false)
      : NLSESolverBase(g, host_u0, host_m, p, sewi ? NLS_INT_SEWI : NLS_INT_SS2) {}
  NLSESolverDevice(const nls::Grid &g, const Scalar *host_u0, const Parameters &p = Parameters()) // gen-1
      : NLSESolverBase(g, host_u0, nullptr, p) {}
  // reference signatures (nlse_dev.hpp:78-82; gen-1 device/nlse_solver_dev.hpp:53-55)
  template <class SpMat, class = decltype(std::declval<const SpMat &>().outerIndexPtr())>
  NLSESolverDevice(const SpMat &L, const Scalar *host_u0, const double *host_m, const bool &is_3d,
                   const Parameters &p = Parameters(), bool sewi = false)
      : NLSESolverBase(L, is_3d ? 3 : 2, host_u0, host_m, p, sewi ? NLS_INT_SEWI : NLS_INT_SS2) {}
  template <class SpMat, class = decltype(std::declval<const SpMat &>().outerIndexPtr())>
  NLSESolverDevice(const SpMat &L, const Scalar *host_u0, const Parameters &p = Parameters())
      : NLSESolverBase(L, 0, host_u0, nullptr, p, NLS_INT_SS2) {}
};

class NLSECubicQuinticSolverDevice : public NLSESolverBase<NLS_EQ_NLSE_CQ> { // nlse_cubic_quintic_dev.hpp:15-127
public:
  NLSECubicQuinticSolverDevice(const nls::Grid &g, const Scalar *host_u0, const double *host_m, const Parameters &p)
      : NLSESolverBase(g, host_u0, host_m, p) {}
  template <class SpMat, class = decltype(std::declval<const SpMat &>().outerIndexPtr())>
  NLSECubicQuinticSolverDevice(const SpMat &L, const Scalar *host_u0, const double *host_m, const Parameters &p)
      : NLSESolverBase(L, 0, host_u0, host_m, p, NLS_INT_SS2) {}
};

class NLSESaturatingSolverDevice : public NLSESolverBase<NLS_EQ_NLSE_SAT> { // nlse_saturating_dev.hpp:15-125
public:
  NLSESaturatingSolverDevice(const nls::Grid &g, const Scalar *host_u0, const double *host_m, const Parameters &p)
      : NLSESolverBase(g, host_u0, host_m, p) {}
  template <class SpMat, class = decltype(std::declval<const SpMat &>().outerIndexPtr())>
  NLSESaturatingSolverDevice(const SpMat &L, const Scalar *host_u0, const double *host_m, const Parameters &p)
      : NLSESolverBase(L, 0, host_u0, host_m, p, NLS_INT_SS2) {}
};

// real wave equations: Gautschi step() and Stoermer-Verlet step_sv() (kg_dev.hpp:93-110)
template <int EQ> class RealSolverBase {
public:
  RealSolverBase(const nls::Grid &g, const double *h_m, const double *host_u0, const double *host_v0, double dt,
                 const Parameters &p, bool sv)
      : grid_(g), h_(g, EQ, sv ? NLS_INT_SV : NLS_INT_GAUTSCHI, p.krylov_dim, dt), p_(p), n_(g.n()), dt_(dt) {
    init(h_m, host_u0, host_v0);
  }
  // reference signature: DEVICE CSR arrays (sg_single_dev.hpp:28-31, kg_dev.hpp:28-31); they stay owned by the caller
  RealSolverBase(const int *d_row_ptr, const int *d_col_ind, const double *d_values, const double *h_m, uint32_t n,
                 uint32_t nnz, const double *host_u0, const double *host_v0, double dt, int dim_hint, const Parameters &p, bool sv)
      : grid_(nls::grid_from_csr(view(d_row_ptr, d_col_ind, d_values, n, nnz), dim_hint)),
        h_(grid_, EQ, sv ? NLS_INT_SV : NLS_INT_GAUTSCHI, p.krylov_dim, dt), p_(p), n_(n), dt_(dt) {
    nls::verify_csr(h_, view(d_row_ptr, d_col_ind, d_values, n, nnz));
    init(h_m, host_u0, host_v0);
  }
  ~RealSolverBase() {
    try { flush(); h_.wait(); } catch (...) {}
  }
  void step() { ++pending_; }
  void step_sv() { ++pending_; } // integrator fixed at construction
  void step(double /*dt*/, uint32_t step_number) { // gen-1 form (device/sg_solver_dev.hpp:201-225): snapshots inside step
    ++pending_;
    if (p_.snapshot_freq && step_number % p_.snapshot_freq == 0) store_snapshot(step_number / p_.snapshot_freq);
  }
  void apply_bc() {}
  void store_snapshot(uint32_t idx) { // sg_single_dev.hpp:96-100 (+ velocity, common_kernels.cuh:6-13)
    if (idx >= p_.num_snapshots) return;
    flush();
    h_.get_async(NLS_FIELD_U, u_traj_.data() + (size_t)idx * n_, n_ * 8);
    h_.get_async(NLS_FIELD_V, v_traj_.data() + (size_t)idx * n_, n_ * 8);
  }
  void transfer_snapshots(double *dst_u, double *dst_v = nullptr) {
    flush();
    h_.wait();
    std::memcpy(dst_u, u_traj_.data(), u_traj_.size() * 8);
    if (dst_v) std::memcpy(dst_v, v_traj_.data(), v_traj_.size() * 8);
  }
  void transfer_snapshots(double *dst, char which) { // gen-1 form: 'u' or 'v'
    flush();
    h_.wait();
    std::memcpy(dst, which == 'v' ? v_traj_.data() : u_traj_.data(), u_traj_.size() * 8);
  }
  nls::Handle &handle() { flush(); return h_; }
  const nls::Grid &grid() const { return grid_; }

protected:
  static nls::CsrView view(const int *rp, const int *ci, const double *val, uint32_t n, uint32_t nnz) {
    nls::CsrView v;
    v.row_ptr = rp; v.col_ind = ci; v.values = val; v.n = n; v.nnz = nnz;
    v.values_complex = false;
    v.on_device = nls_pointer_is_device(rp) != 0;
    return v;
  }
  void init(const double *h_m, const double *host_u0, const double *host_v0) {
    std::vector<double> ones;
    if (!h_m) { ones.assign(n_, 1.0); h_m = ones.data(); }
    h_.set(NLS_FIELD_M, h_m, n_ * 8);
    h_.set(NLS_FIELD_U, host_u0, n_ * 8);
    h_.set(NLS_FIELD_V, host_v0, n_ * 8); // u_past = u0 - dt*v0 (sg_single_driver.cpp:70)
    u_traj_.resize((size_t)p_.num_snapshots * n_);
    v_traj_.resize((size_t)p_.num_snapshots * n_);
    if (p_.num_snapshots) {
      std::memcpy(u_traj_.data(), host_u0, n_ * 8);
      std::memcpy(v_traj_.data(), host_v0, n_ * 8);
    }
  }
  void flush() {
    if (pending_) h_.step(dt_, pending_);
    pending_ = 0;
  }
  nls::Grid grid_;
  nls::Handle h_;
  Parameters p_;
  size_t n_;
  double dt_;
  uint32_t pending_ = 0;
  nls::Pinned<double> u_traj_, v_traj_;
};

#define NLS_REAL_SOLVER(NAME, EQ)                                                                          \
  class NAME : public RealSolverBase<EQ> {                                                                 \
  public:                                                                                                  \
    NAME(const nls::Grid &g, const double *h_m, const double *host_u0, const double *host_v0, double dt,   \
         const Parameters &p = Parameters(), bool stoermer_verlet = false)                                 \
        : RealSolverBase(g, h_m, host_u0, host_v0, dt, p, stoermer_verlet) {}                              \
    /* sg_single_dev.hpp:28-31 */                                                                          \
    NAME(const int *d_row_ptr, const int *d_col_ind, const double *d_values, const double *h_m, uint32_t n, \
         uint32_t nnz, const double *host_u0, const double *host_v0, double dt,                            \
         const Parameters &p = Parameters(), bool stoermer_verlet = false)                                 \
        : RealSolverBase(d_row_ptr, d_col_ind, d_values, h_m, n, nnz, host_u0, host_v0, dt, 0, p,          \
                         stoermer_verlet) {}                                                               \
    /* kg_dev.hpp:28-31 */                                                                                 \
    NAME(const int *d_row_ptr, const int *d_col_ind, const double *d_values, const double *h_m, uint32_t n, \
         uint32_t nnz, const double *host_u0, const double *host_v0, double dt, const bool is_3d,          \
         const Parameters &p, bool stoermer_verlet = false)                                                \
        : RealSolverBase(d_row_ptr, d_col_ind, d_values, h_m, n, nnz, host_u0, host_v0, dt, is_3d ? 3 : 2, \
                         p, stoermer_verlet) {}                                                            \
  };
NLS_REAL_SOLVER(SGESolverDevice, NLS_EQ_SGE)                 // sg_single_dev.hpp:15-121
NLS_REAL_SOLVER(KGESolverDevice, NLS_EQ_KGE)                 // kg_dev.hpp:15-153
NLS_REAL_SOLVER(Phi4SolverDevice, NLS_EQ_PHI4)               // phi4_dev.hpp
NLS_REAL_SOLVER(SGEDoubleSolverDevice, NLS_EQ_SGE_DOUBLE)    // sg_double_dev.hpp
NLS_REAL_SOLVER(SGEHyperbolicSolverDevice, NLS_EQ_SGE_SINH)  // sg_hyperbolic_dev.hpp
#undef NLS_REAL_SOLVER

} // namespace device

// -------------------------------------------------------------------------------------------
// matrix-function applicators and the SpMV type (device/matfunc_complex.hpp:64-230,
// device/matfunc_real.hpp:84-269, device/spmv.hpp:7-73).  apply()/multiply() take DEVICE pointers like the
// reference (stream-ordered on the handle's stream) and, as a convenience, host pointers (blocking).
enum class FunctionType { COS_SQRT, SINC2_SQRT, ID_SQRT }; // device/matfunc_real.hpp:86

class MatrixFunctionApplicatorComplex {
public:
  MatrixFunctionApplicatorComplex(const nls::Grid &g, uint32_t m)
      : grid_(g), h_(g, NLS_EQ_NLSE_CUBIC, NLS_INT_SS2, m), n_(g.n()) {}
  // device/matfunc_complex.hpp:71-73: (A, n, m, nnz, block_size)
  template <class SpMat, class = decltype(std::declval<const SpMat &>().outerIndexPtr())>
  MatrixFunctionApplicatorComplex(const SpMat &A, uint32_t /*n*/, uint32_t m, uint32_t /*nnz*/ = 0, uint32_t /*block_size*/ = 256)
      : grid_(nls::grid_from_csr(nls::csr_view_of(A))), h_(grid_, NLS_EQ_NLSE_CUBIC, NLS_INT_SS2, m), n_(grid_.n()) {
    nls::verify_csr(h_, nls::csr_view_of(A));
  }
  // result = f(dt * L) input; func = "exp" | "sinc" (gen-2 matfunc_complex.hpp:254-256).  The reference's pointer
  // type is thrust::complex<double>*, layout-compatible with std::complex<double>*.
  template <class Cplx>
  void apply(Cplx *result, const Cplx *input, std::complex<double> dt, const std::string &func = "exp") {
    static_assert(sizeof(Cplx) == 16, "complex128 pointers expected");
    const int fn = func == "sinc" ? NLS_FN_SINC : NLS_FN_EXP;
    if (func != "exp" && func != "sinc") throw std::runtime_error("unknown matrix function " + func);
    if (nls_pointer_is_device(input))
      nls::check(nls_matfunc_dev(h_.raw(), fn, dt.real(), dt.imag(), input, result), "nls_matfunc_dev");
    else
      nls::check(nls_matfunc(h_.raw(), fn, dt.real(), dt.imag(), input, result), "nls_matfunc");
  }
  void sync() { h_.sync(); }
  nls::Handle &handle() { return h_; }

private:
  nls::Grid grid_;
  nls::Handle h_;
  size_t n_;
};

class MatrixFunctionApplicatorReal {
public:
  MatrixFunctionApplicatorReal(const nls::Grid &g, uint32_t m)
      : grid_(g), h_(g, NLS_EQ_SGE, NLS_INT_GAUTSCHI, m), n_(g.n()) {}
  // device/matfunc_real.hpp:111-114: (d_row_ptr, d_col_ind, d_values, n, m, nnz, block_size)
  MatrixFunctionApplicatorReal(const int *d_row_ptr, const int *d_col_ind, const double *d_values, uint32_t n, uint32_t m,
                               uint32_t nnz, uint32_t /*block_size*/ = 256)
      : grid_(nls::grid_from_csr(view(d_row_ptr, d_col_ind, d_values, n, nnz))), h_(grid_, NLS_EQ_SGE, NLS_INT_GAUTSCHI, m), n_(n) {
    nls::verify_csr(h_, view(d_row_ptr, d_col_ind, d_values, n, nnz));
  }
  void apply(double *result, const double *input, double t, FunctionType type) {
    const int fn = type == FunctionType::COS_SQRT ? NLS_FN_COS_SQRT
                   : type == FunctionType::SINC2_SQRT ? NLS_FN_SINC2_SQRT : NLS_FN_ID_SQRT;
    if (nls_pointer_is_device(input)) nls::check(nls_matfunc_dev(h_.raw(), fn, t, 0.0, input, result), "nls_matfunc_dev");
    else nls::check(nls_matfunc(h_.raw(), fn, t, 0.0, input, result), "nls_matfunc");
  }
  void sync() { h_.sync(); }
  nls::Handle &handle() { return h_; }

private:
  static nls::CsrView view(const int *rp, const int *ci, const double *val, uint32_t n, uint32_t nnz) {
    nls::CsrView v;
    v.row_ptr = rp; v.col_ind = ci; v.values = val; v.n = n; v.nnz = nnz;
    v.on_device = nls_pointer_is_device(rp) != 0;
    return v;
  }
  nls::Grid grid_;
  nls::Handle h_;
  size_t n_;
};

template <class T> class DeviceSpMV { // y = L x  (device/spmv.hpp:7-73)
public:
  explicit DeviceSpMV(const nls::Grid &g) : grid_(g), h_(g, eq(), integ(), 2) {}
  // device/spmv.hpp:30-52: (d_offsets, d_columns, d_values, n, nnz); the arrays stay owned by the caller
  DeviceSpMV(const int *d_offsets, const int *d_columns, const T *d_values, uint32_t n, uint32_t nnz)
      : grid_(nls::grid_from_csr(view(d_offsets, d_columns, d_values, n, nnz))), h_(grid_, eq(), integ(), 2) {
    nls::verify_csr(h_, view(d_offsets, d_columns, d_values, n, nnz));
  }
  // device pointers (the reference's contract): stream-ordered; host pointers: blocking
  void multiply(const T *x, T *y) {
    if (nls_pointer_is_device(x)) nls::check(nls_stencil_apply_dev(h_.raw(), x, y), "nls_stencil_apply_dev");
    else nls::check(nls_stencil_apply(h_.raw(), x, y), "nls_stencil_apply");
  }
  void sync() { h_.sync(); }
  nls::Handle &handle() { return h_; }

private:
  static int eq() { return sizeof(T) == 16 ? NLS_EQ_NLSE_CUBIC : NLS_EQ_SGE; }
  static int integ() { return sizeof(T) == 16 ? NLS_INT_SS2 : NLS_INT_SV; }
  static nls::CsrView view(const int *rp, const int *ci, const T *val, uint32_t n, uint32_t nnz) {
    nls::CsrView v;
    v.row_ptr = rp; v.col_ind = ci; v.values = val; v.n = n; v.nnz = nnz;
    v.values_complex = sizeof(T) == 16;
    v.on_device = nls_pointer_is_device(rp) != 0;
    return v;
  }
  nls::Grid grid_;
  nls::Handle h_;
};

// gen-1 names live in the global namespace in the reference (device/nlse_solver_dev.hpp:41, device/sg_solver_dev.hpp:92)
using NLSESolverDevice = device::NLSESolverDevice;
using SGESolverDevice = device::SGESolverDevice;
