// ref_harness.cpp -- C entry points around the REFERENCE'S OWN host sources, compiled unmodified from
// /root/reference (paths given with -I by oracle/Makefile) against the mini-Eigen stand-in of
// oracle/ref_shim.  TEST INFRASTRUCTURE ONLY: it exists to check the CPU restatement
// (nls_oracle.cpp) against the reference's source text; outputs go to oracle/_ref/ only.
// Nothing of the reference is copied here: the headers are #included where they lie.
#include <Eigen/Dense>
#include <Eigen/Sparse>

#include <complex>
#include <cstdint>
#include <cstring>
#include <tuple>

#include "laplacians.hpp"            // nlsolvers/common/include
#include "eigen_krylov_complex.hpp"  // nlsolvers/host/include
#include "eigen_krylov_real.hpp"
#include "boundaries.hpp"
#include "boundaries_3d.hpp"
#include "nlse_cubic_solver.hpp"
#include "nlse_cubic_quintic_solver.hpp"
#include "nlse_saturating_solver.hpp"
#include "nlse_cubic_gautschi_solver.hpp"
#include "nlse_cubic_gautschi_solver_3d.hpp"
#include "nlse_cubic_quintic_gautschi_solver.hpp"
#include "nlse_saturating_gautschi_solver.hpp"
#include "sg_single_solver.hpp"
#undef SG_SINGLE_SOLVER_HPP // sg_single_sv_solver.hpp re-uses the include guard of sg_single_solver.hpp
#include "sg_single_sv_solver.hpp"
#include "kg_solver.hpp"
#include "kg_sv_solver.hpp"
#include "phi4_solver.hpp"
#include "sg_double_solver.hpp"
#include "sg_hyperbolic_solver.hpp"
#include "sg_hyperbolic_sv_solver.hpp"
#include "kg_solver_3d.hpp"
#include "kg_sv_solver_3d.hpp"
#undef NLSE_CUBIC_SOLVER_HPP // nlse_cubic_solver_3d.hpp re-uses the include guard of nlse_cubic_solver.hpp
#include "nlse_cubic_solver_3d.hpp"

using cplx = std::complex<double>;

struct ref_ctx {
  uint32_t nx, ny, nz;
  Eigen::SparseMatrix<double> Lr;
  Eigen::SparseMatrix<cplx> Lc;
};

extern "C" {

// kind: 0 build_laplacian_noflux, 1 build_anisotropic_laplacian_noflux, 2 build_laplacian_noflux_3d,
// 3 build_anisotropic_laplacian_noflux_3d; called exactly as the drivers do (nx-2, ny-2[, nz-2]) and
// then scaled by `sign` (the KGE/phi4 drivers pass -L, kg_driver.cpp:93)
ref_ctx *ref_create(int kind, uint32_t nx, uint32_t ny, uint32_t nz, double dx, double dy, double dz, const double *c,
                    double sign) {
  auto *h = new ref_ctx{nx, ny, nz, {}, {}};
  const long n = (long)nx * ny * nz;
  Eigen::VectorX<double> cr(n);
  Eigen::VectorX<cplx> cc(n);
  for (long i = 0; i < n; ++i) { cr[i] = c ? c[i] : 1.0; cc[i] = cr[i]; }
  switch (kind) {
  case 0:
    h->Lr = build_laplacian_noflux<double>(nx - 2, ny - 2, dx, dy);
    h->Lc = build_laplacian_noflux<cplx>(nx - 2, ny - 2, cplx(dx), cplx(dy));
    break;
  case 1:
    h->Lr = build_anisotropic_laplacian_noflux<double>(nx - 2, ny - 2, dx, dy, cr);
    h->Lc = build_anisotropic_laplacian_noflux<cplx>(nx - 2, ny - 2, cplx(dx), cplx(dy), cc);
    break;
  case 2:
    h->Lr = build_laplacian_noflux_3d<double>(nx - 2, ny - 2, nz - 2, dx, dy, dz);
    h->Lc = build_laplacian_noflux_3d<cplx>(nx - 2, ny - 2, nz - 2, cplx(dx), cplx(dy), cplx(dz));
    break;
  default:
    h->Lr = build_anisotropic_laplacian_noflux_3d<double>(nx - 2, ny - 2, nz - 2, dx, dy, dz, cr);
    h->Lc = build_anisotropic_laplacian_noflux_3d<cplx>(nx - 2, ny - 2, nz - 2, cplx(dx), cplx(dy), cplx(dz), cc);
  }
  if (sign != 1.0) { h->Lr *= sign; h->Lc *= cplx(sign); }
  return h;
}
void ref_destroy(ref_ctx *h) { delete h; }

// the assembled matrix in compressed-column form (symmetric, so equally its CSR)
long ref_nnz(ref_ctx *h) { return h->Lr.nonZeros(); }
void ref_matrix(ref_ctx *h, int64_t *outer, int32_t *inner, double *val) {
  for (size_t i = 0; i < h->Lr.outer.size(); ++i) outer[i] = h->Lr.outer[i];
  for (size_t i = 0; i < h->Lr.inner.size(); ++i) { inner[i] = (int32_t)h->Lr.inner[i]; val[i] = h->Lr.val[i]; }
}

void ref_apply(ref_ctx *h, int is_cplx, const void *x, void *y) {
  const long n = h->Lr.rows();
  if (is_cplx) {
    Eigen::VectorX<cplx> v(n);
    std::memcpy(v.data(), x, n * 16);
    Eigen::VectorX<cplx> r = h->Lc * v;
    std::memcpy(y, r.data(), n * 16);
  } else {
    Eigen::VectorX<double> v(n);
    std::memcpy(v.data(), x, n * 8);
    Eigen::VectorX<double> r = h->Lr * v;
    std::memcpy(y, r.data(), n * 8);
  }
}

// lanczos_L_complex / lanczos_L: V (n x m column major), T (m x m column major), beta
void ref_lanczos(ref_ctx *h, int is_cplx, const void *u, uint32_t m, void *V, void *T, double *beta) {
  const long n = h->Lr.rows();
  if (is_cplx) {
    Eigen::VectorX<cplx> v(n);
    std::memcpy(v.data(), u, n * 16);
    auto [Vm, Tm, b] = lanczos_L_complex<cplx>(h->Lc, v, m);
    std::memcpy(V, Vm.data(), (size_t)n * m * 16);
    std::memcpy(T, Tm.data(), (size_t)m * m * 16);
    *beta = b.real();
  } else {
    Eigen::VectorX<double> v(n);
    std::memcpy(v.data(), u, n * 8);
    auto [Vm, Tm, b] = lanczos_L<double>(h->Lr, v, m);
    std::memcpy(V, Vm.data(), (size_t)n * m * 8);
    std::memcpy(T, Tm.data(), (size_t)m * m * 8);
    *beta = b;
  }
}

// fn numbering of include/nls.h (NLS_FN_*)
void ref_matfunc(ref_ctx *h, int fn, double t_re, double t_im, const void *in, void *out, uint32_t m) {
  const long n = h->Lr.rows();
  if (fn <= 1) {
    Eigen::VectorX<cplx> v(n);
    std::memcpy(v.data(), in, n * 16);
    Eigen::VectorX<cplx> r = fn == 0 ? expm_multiply<cplx>(h->Lc, v, cplx(t_re, t_im), m)
                                     : sincm_multiply<cplx>(h->Lc, v, cplx(t_re, t_im), m);
    std::memcpy(out, r.data(), n * 16);
    return;
  }
  Eigen::VectorX<double> v(n), r;
  std::memcpy(v.data(), in, n * 8);
  switch (fn) {
  case 2: r = cos_sqrt_multiply<double>(h->Lr, v, t_re, m); break;
  case 3: r = sinc2_sqrt_multiply<double>(h->Lr, v, t_re, m); break;
  case 4: r = id_sqrt_multiply<double>(h->Lr, v, t_re, m); break;
  case 5: r = sinc2_sqrt_half<double>(h->Lr, v, t_re, m); break;
  default: r = mod_cosine_multiply<double>(h->Lr, v, t_re, m);
  }
  std::memcpy(out, r.data(), n * 8);
}

static void bc_c(ref_ctx *h, Eigen::VectorX<cplx> &u) {
  if (h->nz > 1) neumann_bc_no_velocity_3d<cplx>(u, h->nx, h->ny, h->nz);
  else neumann_bc_no_velocity<cplx>(u, h->nx, h->ny);
}
static void bc_r(ref_ctx *h, Eigen::VectorX<double> &u) {
  if (h->nz > 1) neumann_bc_no_velocity_3d<double>(u, h->nx, h->ny, h->nz);
  else neumann_bc_no_velocity<double>(u, h->nx, h->ny);
}

// nsteps driver-loop bodies "step + boundary copy" of the NLSE solvers.
// kind: 0 NLSESolver, 1 NLSEQuinticSolver, 2 NLSESaturatingSolver (SS2);
//       3 NLSESolver3d (SS2, nlse_cubic_solver_3d.hpp);
//       10 NLSECubicGautschiSolver (2D host convention), 11 NLSEGautschiSolver3d, 12 cq, 13 saturating (sEWI:
//       u_prev in/out; the caller performs the SS2 start-up step with kind 0..2 like the drivers do)
void ref_nlse_steps(ref_ctx *h, int kind, void *u_io, void *uprev_io, const double *mfield, double dt, double s1, double s2,
                    double kappa, uint32_t nsteps) {
  const long n = h->Lr.rows();
  Eigen::VectorX<cplx> u(n), up(n), buf(n), rho(n), mc(n);
  Eigen::VectorX<double> m(n);
  std::memcpy(u.data(), u_io, n * 16);
  if (uprev_io) std::memcpy(up.data(), uprev_io, n * 16);
  for (long i = 0; i < n; ++i) { m[i] = mfield[i]; mc[i] = mfield[i]; }
  const cplx tau(0.0, dt);
  for (uint32_t s = 0; s < nsteps; ++s) {
    switch (kind) {
    case 0: NLSESolver::step<cplx>(buf, rho, u, h->Lc, m, tau); break;
    case 1: NLSEQuinticSolver::step<cplx>(buf, rho, u, h->Lc, m, tau, s1, s2); break;
    case 2: NLSESaturatingSolver::step<cplx>(buf, rho, u, h->Lc, m, tau, kappa); break;
    case 3: NLSESolver3d::step<cplx>(buf, rho, u, h->Lc, mc, tau); break; // m declared complex there (nlse_cubic_solver_3d.hpp:12)
    case 10: NLSECubicGautschiSolver::step<cplx>(buf, rho, u, up, h->Lc, m, tau); break;
    case 11: NLSEGautschiSolver3d::step<cplx>(buf, rho, u, up, h->Lc, mc, tau); break;
    case 12: NLSECubicQuinticGautschiSolver::step<cplx>(buf, rho, u, up, h->Lc, m, tau, s1, s2); break;
    default: NLSESaturatingGautschiSolver::step<cplx>(buf, rho, u, up, h->Lc, m, tau, kappa);
    }
    bc_c(h, u);
  }
  std::memcpy(u_io, u.data(), n * 16);
  if (uprev_io) std::memcpy(uprev_io, up.data(), n * 16);
}

// kind: 0 SGESolver (Gautschi), 1 SGESolverSV, 2 KGESolver, 3 KGESVSolver, 4 Phi4Solver, 5 SGEDoubleSolver,
//       6 SGEHyperbolicSolver, 7 SGEHyperbolicSVSolver, 8 KGESolver3d, 9 KGESVSolver3d
void ref_real_steps(ref_ctx *h, int kind, double *u_io, double *upast_io, const double *mfield, double dt, uint32_t nsteps) {
  const long n = h->Lr.rows();
  Eigen::VectorX<double> u(n), up(n), buf(n), m(n);
  std::memcpy(u.data(), u_io, n * 8);
  std::memcpy(up.data(), upast_io, n * 8);
  std::memcpy(m.data(), mfield, n * 8);
  for (uint32_t s = 0; s < nsteps; ++s) {
    switch (kind) {
    case 0: SGESolver::step<double>(u, up, buf, h->Lr, m, dt); break;
    case 1: SGESolverSV::step<double>(u, up, buf, h->Lr, m, dt); break;
    case 2: KGESolver::step<double>(u, up, buf, h->Lr, m, dt); break;
    case 3: KGESVSolver::step<double>(u, up, buf, h->Lr, m, dt); break;
    case 4: Phi4Solver::step<double>(u, up, buf, h->Lr, m, dt); break;
    case 5: SGEDoubleSolver::step<double>(u, up, buf, h->Lr, m, dt); break;
    case 6: SGEHyperbolicSolver::step<double>(u, up, buf, h->Lr, m, dt); break;
    case 7: SGEHyperbolicSVSolver::step<double>(u, up, buf, h->Lr, m, dt); break;
    case 8: KGESolver3d::step<double>(u, up, buf, h->Lr, m, dt); break;
    default: KGESVSolver3d::step<double>(u, up, buf, h->Lr, m, dt);
    }
    bc_r(h, u);
  }
  std::memcpy(u_io, u.data(), n * 8);
  std::memcpy(upast_io, up.data(), n * 8);
}

} // extern "C"
