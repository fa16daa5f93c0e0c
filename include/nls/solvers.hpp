// solvers.hpp -- the reference's C++ solver-class surface (SURVEY.md 8b "B2") over the C ABI
// of include/nls.h.  Header-only host C++17, no CUDA, no Eigen: link with libnls_b200.so.
//
// The reference classes take an Eigen::SparseMatrix only to ship a CSR to the GPU
// (device/nlse_solver_dev.hpp:53-55, nlsolvers/device/include/nlse_dev.hpp:78-82,
// sg_single_dev.hpp:28-31).  The engine is matrix-free, so the drop-in constructors take the
// grid geometry (+ optional c field) instead: `nls::Grid` replaces the matrix argument, every
// other argument keeps its position and meaning.  Errors throw std::runtime_error like the
// reference (nlse_solver_dev.hpp:77-81).  Differences that a caller can observe:
//   * step() includes the no-flux boundary copy (the drivers always call apply_bc() right
//     after step(); apply_bc() is kept as a no-op),
//   * snapshots are collected in host memory owned by the object (transfer_snapshots copies
//     them out) instead of a device-resident trajectory.
#pragma once
#include "../nls.h"

#include <complex>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace nls {

struct Grid {
  uint32_t nx = 0, ny = 0, nz = 1;
  double dx = 0, dy = 0, dz = 0;
  const double *c = nullptr; // host c(x) field => anisotropic Laplacian (laplacians.hpp:54-103,:158-218)
  double op_sign = 1.0;      // -1: the drivers that pass -L (kg_driver.cpp:93)
  size_t n() const { return (size_t)nx * ny * nz; }
};

inline void check(int rc, const char *what) {
  if (rc != NLS_OK) throw std::runtime_error(std::string(what) + ": " + nls_last_error());
}

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
    if (g.c) set(NLS_FIELD_C, g.c, n_ * sizeof(double));
  }
  ~Handle() { nls_destroy(h_); }
  Handle(const Handle &) = delete;
  Handle &operator=(const Handle &) = delete;

  void set(int field, const void *src, size_t bytes) { check(nls_set_field(h_, field, src, bytes), "nls_set_field"); }
  void get(int field, void *dst, size_t bytes) { check(nls_get_field(h_, field, dst, bytes), "nls_get_field"); }
  // field resampled on the device to tnx x tny [x tnz] points per trajectory (downsampling.py:68-104)
  void get_resampled(int field, uint32_t tnx, uint32_t tny, uint32_t tnz, void *dst, size_t bytes) {
    check(nls_get_field_resampled(h_, field, tnx, tny, tnz, dst, bytes), "nls_get_field_resampled");
  }
  // NLS_RED_MASS / NLS_RED_ENERGY / NLS_RED_NONFINITE, one value per trajectory
  std::vector<double> reduce(int what) {
    std::vector<double> out(batch_);
    check(nls_reduce(h_, what, out.data()), "nls_reduce");
    return out;
  }
  void step(double dt, uint32_t nsteps = 1) { check(nls_step(h_, dt, nsteps), "nls_step"); }
  void sync() { check(nls_sync(h_), "nls_sync"); }
  nls_handle raw() const { return h_; }
  size_t n() const { return n_; }
  size_t state_bytes() const { return n_ * batch_ * (cplx_ ? 16 : 8); }

private:
  nls_handle h_ = nullptr;
  size_t n_ = 0, batch_ = 1;
  bool cplx_ = false;
};

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

// common NLSE machinery: SS2 step, snapshot bookkeeping of nlse_solver_dev.hpp:94-124
template <int EQ> class NLSESolverBase {
public:
  using Scalar = std::complex<double>;
  NLSESolverBase(const nls::Grid &g, const Scalar *host_u0, const double *host_m, const Parameters &p,
                 int integrator = NLS_INT_SS2)
      : h_(g, EQ, integrator, p.krylov_dim, 0.0, p.sigma1, p.sigma2, p.kappa), p_(p), n_(g.n()) {
    h_.set(NLS_FIELD_U, host_u0, n_ * sizeof(Scalar));
    std::vector<double> ones;
    if (!host_m) { ones.assign(n_, 1.0); host_m = ones.data(); } // nlse_cubic_quintic_driver_dev.cpp:83-85
    h_.set(NLS_FIELD_M, host_m, n_ * sizeof(double));
    traj_.resize((size_t)p.num_snapshots * n_);
    if (p.num_snapshots) { std::memcpy(traj_.data(), host_u0, n_ * sizeof(Scalar)); current_ = 1; }
  }
  // tau = i*dt (nlse_cubic_driver.cpp:50); only its imaginary part is meaningful
  void step(const Scalar tau, const uint32_t step_number) {
    h_.step(tau.imag(), 1);
    if (p_.snapshot_freq && step_number % p_.snapshot_freq == 0 && current_ < p_.num_snapshots)
      h_.get(NLS_FIELD_U, traj_.data() + (size_t)current_++ * n_, n_ * sizeof(Scalar));
  }
  // gen-2 signature: online snapshot into caller memory (nlse_dev.hpp:187, :323-334)
  void step(const Scalar tau, const uint32_t step_number, Scalar *host_dst) {
    h_.step(tau.imag(), 1);
    if (host_dst && p_.snapshot_freq && step_number % p_.snapshot_freq == 0) {
      const uint32_t idx = step_number / p_.snapshot_freq;
      if (idx < p_.num_snapshots) h_.get(NLS_FIELD_U, host_dst + (size_t)idx * n_, n_ * sizeof(Scalar));
    }
  }
  // sEWI step (nlse_dev.hpp:205-238): construct the solver with sewi = true; the first call is the SS2
  // start-up step with u- = u0, exactly as the reference's `if (step_number == 1)` branch
  void step_sewi(const Scalar tau, const uint32_t step_number, Scalar *host_dst) { step(tau, step_number, host_dst); }
  void apply_bc() {} // fused into step()
  void transfer_snapshots(Scalar *dst) { std::memcpy(dst, traj_.data(), traj_.size() * sizeof(Scalar)); }
  void current_state(Scalar *dst) { h_.get(NLS_FIELD_U, dst, n_ * sizeof(Scalar)); }
  nls::Handle &handle() { return h_; }

protected:
  nls::Handle h_;
  Parameters p_;
  size_t n_;
  uint32_t current_ = 0;
  std::vector<Scalar> traj_;
};

class NLSESolverDevice : public NLSESolverBase<NLS_EQ_NLSE_CUBIC> { // nlse_dev.hpp:66-361
public:
  NLSESolverDevice(const nls::Grid &g, const Scalar *host_u0, const double *host_m, const bool & /*is_3d*/,
                   const Parameters &p = Parameters(), bool sewi = false)
      : NLSESolverBase(g, host_u0, host_m, p, sewi ? NLS_INT_SEWI : NLS_INT_SS2) {}
  NLSESolverDevice(const nls::Grid &g, const Scalar *host_u0, const Parameters &p = Parameters()) // gen-1
      : NLSESolverBase(g, host_u0, nullptr, p) {}
};

class NLSECubicQuinticSolverDevice : public NLSESolverBase<NLS_EQ_NLSE_CQ> { // nlse_cubic_quintic_dev.hpp:15-127
public:
  NLSECubicQuinticSolverDevice(const nls::Grid &g, const Scalar *host_u0, const double *host_m, const Parameters &p)
      : NLSESolverBase(g, host_u0, host_m, p) {}
};

class NLSESaturatingSolverDevice : public NLSESolverBase<NLS_EQ_NLSE_SAT> { // nlse_saturating_dev.hpp:15-125
public:
  NLSESaturatingSolverDevice(const nls::Grid &g, const Scalar *host_u0, const double *host_m, const Parameters &p)
      : NLSESolverBase(g, host_u0, host_m, p) {}
};

// real wave equations: Gautschi step() and Stoermer-Verlet step_sv() (kg_dev.hpp:93-110)
template <int EQ> class RealSolverBase {
public:
  RealSolverBase(const nls::Grid &g, const double *h_m, const double *host_u0, const double *host_v0, double dt,
                 const Parameters &p, bool sv)
      : h_(g, EQ, sv ? NLS_INT_SV : NLS_INT_GAUTSCHI, p.krylov_dim, dt), p_(p), n_(g.n()), dt_(dt) {
    std::vector<double> ones;
    if (!h_m) { ones.assign(n_, 1.0); h_m = ones.data(); }
    h_.set(NLS_FIELD_M, h_m, n_ * 8);
    h_.set(NLS_FIELD_U, host_u0, n_ * 8);
    h_.set(NLS_FIELD_V, host_v0, n_ * 8); // u_past = u0 - dt*v0 (sg_single_driver.cpp:70)
    u_traj_.resize((size_t)p.num_snapshots * n_);
    v_traj_.resize((size_t)p.num_snapshots * n_);
    if (p.num_snapshots) {
      std::memcpy(u_traj_.data(), host_u0, n_ * 8);
      std::memcpy(v_traj_.data(), host_v0, n_ * 8);
    }
  }
  void step() { h_.step(dt_, 1); }
  void step_sv() { h_.step(dt_, 1); } // integrator fixed at construction
  void apply_bc() {}
  void store_snapshot(uint32_t idx) { // sg_single_dev.hpp:96-100 (+ velocity, common_kernels.cuh:6-13)
    if (idx >= p_.num_snapshots) return;
    h_.get(NLS_FIELD_U, u_traj_.data() + (size_t)idx * n_, n_ * 8);
    h_.get(NLS_FIELD_V, v_traj_.data() + (size_t)idx * n_, n_ * 8);
  }
  void transfer_snapshots(double *dst_u, double *dst_v = nullptr) {
    std::memcpy(dst_u, u_traj_.data(), u_traj_.size() * 8);
    if (dst_v) std::memcpy(dst_v, v_traj_.data(), v_traj_.size() * 8);
  }
  nls::Handle &handle() { return h_; }

protected:
  nls::Handle h_;
  Parameters p_;
  size_t n_;
  double dt_;
  std::vector<double> u_traj_, v_traj_;
};

#define NLS_REAL_SOLVER(NAME, EQ)                                                                          \
  class NAME : public RealSolverBase<EQ> {                                                                 \
  public:                                                                                                  \
    NAME(const nls::Grid &g, const double *h_m, const double *host_u0, const double *host_v0, double dt,   \
         const Parameters &p = Parameters(), bool stoermer_verlet = false)                                 \
        : RealSolverBase(g, h_m, host_u0, host_v0, dt, p, stoermer_verlet) {}                              \
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
// device/matfunc_real.hpp:84-269, device/spmv.hpp:7-73).  Host pointers in, host pointers out.
enum class FunctionType { COS_SQRT, SINC2_SQRT, ID_SQRT }; // device/matfunc_real.hpp:86

class MatrixFunctionApplicatorComplex {
public:
  MatrixFunctionApplicatorComplex(const nls::Grid &g, uint32_t m)
      : h_(g, NLS_EQ_NLSE_CUBIC, NLS_INT_SS2, m), n_(g.n()) {}
  // result = f(dt * L) input; func = "exp" | "sinc" (gen-2 matfunc_complex.hpp:254-256)
  void apply(std::complex<double> *result, const std::complex<double> *input, std::complex<double> dt,
             const std::string &func = "exp") {
    const int fn = func == "sinc" ? NLS_FN_SINC : NLS_FN_EXP;
    if (func != "exp" && func != "sinc") throw std::runtime_error("unknown matrix function " + func);
    nls::check(nls_matfunc(h_.raw(), fn, dt.real(), dt.imag(), input, result), "nls_matfunc");
  }

private:
  nls::Handle h_;
  size_t n_;
};

class MatrixFunctionApplicatorReal {
public:
  MatrixFunctionApplicatorReal(const nls::Grid &g, uint32_t m)
      : h_(g, NLS_EQ_SGE, NLS_INT_GAUTSCHI, m), n_(g.n()) {}
  void apply(double *result, const double *input, double t, FunctionType type) {
    const int fn = type == FunctionType::COS_SQRT ? NLS_FN_COS_SQRT
                   : type == FunctionType::SINC2_SQRT ? NLS_FN_SINC2_SQRT : NLS_FN_ID_SQRT;
    nls::check(nls_matfunc(h_.raw(), fn, t, 0.0, input, result), "nls_matfunc");
  }

private:
  nls::Handle h_;
  size_t n_;
};

template <class T> class DeviceSpMV { // y = L x  (device/spmv.hpp:65-72); host arrays
public:
  explicit DeviceSpMV(const nls::Grid &g)
      : h_(g, sizeof(T) == 16 ? NLS_EQ_NLSE_CUBIC : NLS_EQ_SGE, sizeof(T) == 16 ? NLS_INT_SS2 : NLS_INT_SV, 2) {}
  void multiply(const T *x, T *y) { nls::check(nls_stencil_apply(h_.raw(), x, y), "nls_stencil_apply"); }

private:
  nls::Handle h_;
};
