// nls_oracle.cpp -- CPU restatement of the reference host algorithm.
// TEST INFRASTRUCTURE ONLY (see nls_oracle.h).  Pinned against the reference's own host
// sources (oracle/_ref, tests/test_reference_pin.py); Eigen's arithmetic itself is substituted
// ("parity unpinned" only there); also anchored by scipy expm_multiply + numpy twin.
//
// Every function cites the reference file:line it restates (paths relative to
// the reference checkout, nlsolvers/ tree is canonical per SURVEY.md 8c).
#include "nls_oracle.h"

#include <algorithm>
#include <cmath>
#include <complex>
#include <cstring>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

using cplx = std::complex<double>;
thread_local std::string g_err;

// ---------------------------------------------------------------------------
// CSR operator, assembled from triplets exactly as the reference does
// (Eigen setFromTriplets sums duplicates; makeCompressed sorts inner indices).
// Weights are real in every builder; the complex build stores them as complex
// (laplacians.hpp:9-52 with Float=complex<double>), so one real CSR serves both.
// ---------------------------------------------------------------------------
struct Csr {
  int64_t n = 0;
  std::vector<int64_t> rowptr;
  std::vector<int32_t> col;
  std::vector<double> val;
};

struct Trip {
  int64_t r, c;
  double v;
};

Csr from_triplets(int64_t n, std::vector<Trip> &t, double scale) {
  std::stable_sort(t.begin(), t.end(), [](const Trip &a, const Trip &b) {
    return a.r != b.r ? a.r < b.r : a.c < b.c;
  });
  Csr A;
  A.n = n;
  A.rowptr.assign(n + 1, 0);
  for (size_t k = 0; k < t.size();) {
    size_t e = k;
    double s = 0.0;
    while (e < t.size() && t[e].r == t[k].r && t[e].c == t[k].c) s += t[e++].v;
    A.col.push_back((int32_t)t[k].c);
    A.val.push_back(s * scale); // L *= 1/(dx*dy) after assembly (laplacians.hpp:49)
    A.rowptr[t[k].r + 1]++;
    k = e;
  }
  for (int64_t i = 0; i < n; ++i) A.rowptr[i + 1] += A.rowptr[i];
  return A;
}

// build_laplacian_noflux(nx-2, ny-2, dx, dy): nlsolvers/common/include/laplacians.hpp:9-52.
// Called with the full grid sizes here (the "+2" in the reference undoes the caller's "-2").
// NB the reference indexes everything with (nx+2) only, i.e. assumes nx == ny (:14).
Csr lap2d_uniform(int64_t NX, int64_t NY, double dx, double dy) {
  const int64_t N = NX * NY;
  std::vector<Trip> t;
  t.reserve(5 * N);
  for (int64_t i = 0; i < N; ++i) { // :23-30 diagonal -4, ring (incl. corners) -3
    double v = -4.0;
    if (i < NX || i >= N - NX || i % NX == 0 || i % NX == NX - 1) v = -3.0;
    t.push_back({i, i, v});
  }
  for (int64_t i = 0; i < N - 1; ++i) // :32-37 x neighbours, no wrap
    if ((i + 1) % NX != 0) {
      t.push_back({i, i + 1, 1.0});
      t.push_back({i + 1, i, 1.0});
    }
  for (int64_t i = 0; i < N - NX; ++i) { // :39-42 y neighbours
    t.push_back({i, i + NX, 1.0});
    t.push_back({i + NX, i, 1.0});
  }
  return from_triplets(N, t, 1.0 / (dx * dy));
}

// build_anisotropic_laplacian_noflux: laplacians.hpp:54-103
Csr lap2d_aniso(int64_t NX, int64_t NY, double dx, double dy, const double *c) {
  const int64_t N = NX * NY;
  std::vector<Trip> t;
  t.reserve(5 * N);
  std::vector<double> diag(N, 0.0);
  for (int64_t i = 0; i < N - 1; ++i) // :75-83
    if ((i + 1) % NX != 0) {
      double a = (c[i] + c[i + 1]) / 2.0;
      t.push_back({i, i + 1, a});
      t.push_back({i + 1, i, a});
      diag[i] += a;
      diag[i + 1] += a;
    }
  for (int64_t i = 0; i < N - NX; ++i) { // :85-91
    double a = (c[i] + c[i + NX]) / 2.0;
    t.push_back({i, i + NX, a});
    t.push_back({i + NX, i, a});
    diag[i] += a;
    diag[i + NX] += a;
  }
  for (int64_t i = 0; i < N; ++i) t.push_back({i, i, -diag[i]}); // :93-95
  return from_triplets(N, t, 1.0 / (dx * dy));
}

// build_laplacian_noflux_3d: laplacians.hpp:105-156 (incl. the y-loop without a
// plane-boundary check, :140-143 -> coupling (i,ny-1,k) <-> (i,0,k+1)).
Csr lap3d_uniform(int64_t NX, int64_t NY, int64_t NZ, double dx) {
  const int64_t P = NX * NY, N = P * NZ;
  std::vector<Trip> t;
  t.reserve(7 * N);
  for (int64_t k = 0; k < NZ; ++k)
    for (int64_t j = 0; j < NY; ++j)
      for (int64_t i = 0; i < NX; ++i) { // :119-131
        const int64_t idx = k * P + j * NX + i;
        const bool face = k == 0 || k == NZ - 1 || j == 0 || j == NY - 1 || i == 0 || i == NX - 1;
        t.push_back({idx, idx, face ? -5.0 : -6.0});
      }
  for (int64_t i = 0; i < N - 1; ++i) // :133-138
    if ((i + 1) % NX != 0) {
      t.push_back({i, i + 1, 1.0});
      t.push_back({i + 1, i, 1.0});
    }
  for (int64_t i = 0; i < N - NX; ++i) { // :140-143 (no plane check: reference quirk Q2)
    t.push_back({i, i + NX, 1.0});
    t.push_back({i + NX, i, 1.0});
  }
  for (int64_t i = 0; i < N - P; ++i) { // :145-148
    t.push_back({i, i + P, 1.0});
    t.push_back({i + P, i, 1.0});
  }
  return from_triplets(N, t, 1.0 / (dx * dx)); // :153
}

// build_anisotropic_laplacian_noflux_3d: laplacians.hpp:158-218
Csr lap3d_aniso(int64_t NX, int64_t NY, int64_t NZ, double dx, const double *c) {
  const int64_t P = NX * NY, N = P * NZ;
  std::vector<Trip> t;
  t.reserve(7 * N);
  std::vector<double> diag(N, 0.0);
  auto link = [&](int64_t a, int64_t b) {
    double w = (c[a] + c[b]) / 2.0;
    t.push_back({a, b, w});
    t.push_back({b, a, w});
    diag[a] += w;
    diag[b] += w;
  };
  for (int64_t i = 0; i < N - 1; ++i) // :181-190
    if ((i + 1) % NX != 0) link(i, i + 1);
  for (int64_t i = 0; i < N - NX; ++i) link(i, i + NX); // :192-198 (wrap quirk kept)
  for (int64_t i = 0; i < N - P; ++i) link(i, i + P);   // :200-206
  for (int64_t i = 0; i < N; ++i) t.push_back({i, i, -diag[i]}); // :208-210
  return from_triplets(N, t, 1.0 / (dx * dx));
}

// Eigen SparseMatrix(ColMajor) * vector accumulates, for each output row, the
// products in increasing column order starting from 0; CSR with sorted columns
// reproduces that order.  sgn is the driver-level negation (kg_driver.cpp:93).
template <class S>
void spmv(const Csr &A, double sgn, const S *x, S *y) {
  for (int64_t r = 0; r < A.n; ++r) {
    S acc = S(0);
    for (int64_t k = A.rowptr[r]; k < A.rowptr[r + 1]; ++k) acc += (sgn * A.val[k]) * x[A.col[k]];
    y[r] = acc;
  }
}

// ---------------------------------------------------------------------------
// BLAS-1 helpers (Eigen semantics: dot(a,b) for complex = a^H b).
// ---------------------------------------------------------------------------
inline double conj_s(double a) { return a; }
inline cplx conj_s(cplx a) { return std::conj(a); }
inline double real_s(double a) { return a; }
inline double real_s(cplx a) { return a.real(); }
inline double abs2_s(double a) { return a * a; }
inline double abs2_s(cplx a) { return a.real() * a.real() + a.imag() * a.imag(); }

template <class S>
S adj_dot(const S *a, const S *b, int64_t n) { // a.adjoint() * b
  S s = S(0);
  for (int64_t i = 0; i < n; ++i) s += conj_s(a[i]) * b[i];
  return s;
}
template <class S>
double norm2(const S *a, int64_t n) {
  double s = 0.0;
  for (int64_t i = 0; i < n; ++i) s += abs2_s(a[i]);
  return std::sqrt(s);
}

// lanczos_L_complex: nlsolvers/host/include/eigen_krylov_complex.hpp:13-52
// lanczos_L (real):  nlsolvers/host/include/eigen_krylov_real.hpp:8-55
// V: m columns of n (column after column), alpha[m] = diag(T) (alpha[m-1] stays 0,
// SURVEY Q10), off[j] = T(j+1,j), returns beta = ||u||.
template <class S>
double lanczos(const Csr &A, double sgn, const S *u, int m, int reorth, std::vector<S> &V,
               std::vector<S> &alpha, std::vector<double> &off) {
  const int64_t n = A.n;
  V.assign((size_t)n * m, S(0)); // :18 Zero(n,m)
  alpha.assign(m, S(0));         // :19 Zero(m,m)
  off.assign(m, 0.0);
  std::vector<S> w(n);
  const double beta = norm2(u, n); // :21
  for (int64_t i = 0; i < n; ++i) V[i] = u[i] / beta; // :22
  for (int j = 0; j < m - 1; ++j) {                   // :24
    S *vj = &V[(size_t)j * n];
    spmv(A, sgn, vj, w.data()); // :25
    if (j > 0) {                // :26-27
      const S *vp = &V[(size_t)(j - 1) * n];
      const double b = off[j - 1];
      for (int64_t i = 0; i < n; ++i) w[i] -= b * vp[i];
    }
    // complex: T(j,j) = V_j^H w (:28); real: w.dot(V_j) (eigen_krylov_real.hpp:29)
    const S a = adj_dot(vj, w.data(), n);
    alpha[j] = a;
    for (int64_t i = 0; i < n; ++i) w[i] -= a * vj[i]; // :29
    if (reorth != NLS_REORTH_NONE) // every other mode of nls.h is the reference's full sweep in a different kernel form
      for (int i2 = 0; i2 <= j; ++i2) { // :32-35 MGS sweep
        const S *vi = &V[(size_t)i2 * n];
        const S c = adj_dot(vi, w.data(), n);
        for (int64_t i = 0; i < n; ++i) w[i] -= c * vi[i];
      }
    off[j] = norm2(w.data(), n); // :38-39 (no breakdown guard, :40-44 commented out)
    S *vn = &V[(size_t)(j + 1) * n];
    for (int64_t i = 0; i < n; ++i) vn[i] = w[i] / off[j]; // :46
  }
  return beta;
}

// Symmetric tridiagonal eigen-decomposition (implicit QL, EISPACK tql2 form).
// Stands in for Eigen::SelfAdjointEigenSolver at eigen_krylov_complex.hpp:68 /
// eigen_krylov_real.hpp:71 (third-party arithmetic; result enters only through
// Q f(Lambda) Q^H e1, which is sign/order invariant).  d: diag (in) -> eigenvalues,
// e[0..m-2]: sub-diagonal, Z: m x m row-major eigenvectors in columns.
bool tql2(int m, std::vector<double> &d, std::vector<double> e, std::vector<double> &Z) {
  Z.assign((size_t)m * m, 0.0);
  for (int i = 0; i < m; ++i) Z[(size_t)i * m + i] = 1.0;
  e.resize(m);
  e[m - 1] = 0.0;
  for (int l = 0; l < m; ++l) {
    int iter = 0, mm;
    do {
      for (mm = l; mm < m - 1; ++mm) {
        double dd = std::fabs(d[mm]) + std::fabs(d[mm + 1]);
        if (std::fabs(e[mm]) <= 2.220446049250313e-16 * dd) break;
      }
      if (mm != l) {
        if (iter++ == 60) return false;
        double g = (d[l + 1] - d[l]) / (2.0 * e[l]);
        double r = std::hypot(g, 1.0);
        g = d[mm] - d[l] + e[l] / (g + (g >= 0 ? std::fabs(r) : -std::fabs(r)));
        double s = 1.0, c = 1.0, p = 0.0;
        int i;
        for (i = mm - 1; i >= l; --i) {
          double f = s * e[i], b = c * e[i];
          e[i + 1] = (r = std::hypot(f, g));
          if (r == 0.0) {
            d[i + 1] -= p;
            e[mm] = 0.0;
            break;
          }
          s = f / r;
          c = g / r;
          g = d[i + 1] - p;
          r = (d[i] - g) * s + 2.0 * c * b;
          d[i + 1] = g + (p = s * r);
          g = c * r - b;
          for (int k = 0; k < m; ++k) {
            f = Z[(size_t)k * m + i + 1];
            Z[(size_t)k * m + i + 1] = s * Z[(size_t)k * m + i] + c * f;
            Z[(size_t)k * m + i] = c * Z[(size_t)k * m + i] - s * f;
          }
        }
        if (r == 0.0 && i >= l) continue;
        d[l] -= p;
        e[l] = g;
        e[mm] = 0.0;
      }
    } while (mm != l);
  }
  return true;
}

// scalar function tables: eigen_krylov_complex.hpp:70-76, :89-96;
// eigen_krylov_real.hpp:79-80, :109-115, :142-144, :161-170, :189-204
cplx fn_eval(int fn, cplx t, double lam) {
  switch (fn) {
  case NLS_FN_EXP: return std::exp(t * lam);
  case NLS_FN_SINC: {
    cplx x = t * lam;
    return std::abs(x) < 1e-8 ? cplx(1) : std::sin(x) / x;
  }
  case NLS_FN_COS_SQRT: return std::cos(t.real() * std::sqrt(std::fabs(lam)));
  case NLS_FN_SINC2_SQRT: {
    double x = t.real() * std::sqrt(std::fabs(lam));
    double s = std::fabs(x) < 1e-8 ? 1.0 : std::sin(x) / x;
    return s * s;
  }
  case NLS_FN_ID_SQRT: return t.real() * std::sqrt(std::fabs(lam));
  case NLS_FN_SINC2_HALF: {
    double x = t.real() / 2. * std::sqrt(std::fabs(lam));
    double s = std::fabs(x) < 1e-8 ? 1.0 : std::sin(x) / x;
    return s * s;
  }
  case NLS_FN_MOD_COSINE: {
    double th = t.real() * std::sqrt(std::fabs(lam));
    double v;
    if (std::fabs(th) < 1e-12) v = 1.0;
    else {
      double h = th / 2.0;
      v = std::cos(h) * std::cos(h) * std::sin(th) / th;
    }
    return v * v; // trailing .square() eigen_krylov_real.hpp:203
  }
  }
  return 0;
}

inline double to_s(cplx z, double) { return z.real(); }
inline cplx to_s(cplx z, cplx) { return z; }

// coefficient vector  coef = f(T) e1 = Q f(Lambda) Q^T e1   (then result = beta * V * coef)
template <class S>
bool matfunc_coef(int fn, cplx t, const std::vector<S> &alpha, const std::vector<double> &off,
                  std::vector<S> &coef) {
  const int m = (int)alpha.size();
  std::vector<double> d(m), e(off.begin(), off.end()), Z;
  for (int i = 0; i < m; ++i) d[i] = real_s(alpha[i]); // Hermitian solver: real diagonal
  if (!tql2(m, d, e, Z)) return false;
  coef.assign(m, S(0));
  for (int i = 0; i < m; ++i) {
    cplx s = 0;
    for (int k = 0; k < m; ++k) s += Z[(size_t)i * m + k] * fn_eval(fn, t, d[k]) * Z[k]; // Z[0*m+k]
    coef[i] = to_s(s, S());
  }
  return true;
}

// expm_multiply & friends: beta * V * f(T) * e1
template <class S>
bool matfunc_apply(const Csr &A, double sgn, int fn, cplx t, const S *in, int m, int reorth,
                   S *out) {
  std::vector<S> V, alpha, coef;
  std::vector<double> off;
  const double beta = lanczos(A, sgn, in, m, reorth, V, alpha, off);
  if (!matfunc_coef(fn, t, alpha, off, coef)) return false;
  const int64_t n = A.n;
  for (int64_t i = 0; i < n; ++i) {
    S s = S(0);
    for (int k = 0; k < m; ++k) s += V[(size_t)k * n + i] * coef[k];
    out[i] = beta * s;
  }
  return true;
}

// ---------------------------------------------------------------------------
// Pointwise physics
// ---------------------------------------------------------------------------
// f(u) of the real equations: sg_single_solver.hpp:53-54, kg_solver.hpp:17,
// phi4_solver.hpp:18-19, sg_double_solver.hpp:17-18, sg_hyperbolic_solver.hpp:18
double f_real(int eq, double x) {
  switch (eq) {
  case NLS_EQ_SGE: return std::sin(x);
  case NLS_EQ_KGE: return x * x * x;
  case NLS_EQ_PHI4: return (x - x * x * x);
  case NLS_EQ_SGE_DOUBLE: return std::sin(x) + .6 * std::sin(2. * x);
  case NLS_EQ_SGE_SINH: return std::sinh(x);
  }
  return 0.0;
}

// neumann_bc_no_velocity: nlsolvers/host/include/boundaries.hpp:40-57 (the Eigen Map is
// column-major on the flat array: "row" = x index), 3D: boundaries_3d.hpp:8-32.
template <class S>
void neumann_bc(S *u, int64_t nx, int64_t ny, int64_t nz) {
  if (nz == 1) {
    for (int64_t j = 1; j < ny - 1; ++j) u[j * nx] = u[j * nx + 1];
    for (int64_t j = 1; j < ny - 1; ++j) u[j * nx + nx - 1] = u[j * nx + nx - 2];
    for (int64_t i = 0; i < nx; ++i) u[i] = u[nx + i];
    for (int64_t i = 0; i < nx; ++i) u[(ny - 1) * nx + i] = u[(ny - 2) * nx + i];
    return;
  }
  const int64_t P = nx * ny;
  for (int64_t k = 1; k < nz - 1; ++k)
    for (int64_t j = 1; j < ny - 1; ++j) {
      u[j * nx + k * P] = u[j * nx + 1 + k * P];
      u[j * nx + (nx - 1) + k * P] = u[j * nx + (nx - 2) + k * P];
    }
  for (int64_t k = 1; k < nz - 1; ++k)
    for (int64_t i = 0; i < nx; ++i) {
      u[i + k * P] = u[i + nx + k * P];
      u[i + (ny - 1) * nx + k * P] = u[i + (ny - 2) * nx + k * P];
    }
  for (int64_t j = 0; j < ny; ++j)
    for (int64_t i = 0; i < nx; ++i) {
      u[i + j * nx] = u[i + j * nx + P];
      u[i + j * nx + (nz - 1) * P] = u[i + j * nx + (nz - 2) * P];
    }
}

} // namespace

struct orc_engine {
  nls_desc d;
  int64_t n;
  bool cplx_state;
  Csr A;
  bool built = false;
  double sgn = 1.0;
  std::vector<double> m, c;
  std::vector<cplx> uc, uc_prev;   // batch*n (uc_prev: sEWI u^{n-1})
  std::vector<char> sewi_started;  // per trajectory
  std::vector<double> ur, ur_past; // batch*n

  void build() {
    if (built) return;
    const bool an = d.laplacian == NLS_LAP_ANISO;
    if (an && c.empty()) c.assign(n, 1.0);
    if (d.nz == 1)
      A = an ? lap2d_aniso(d.nx, d.ny, d.dx, d.dy, c.data()) : lap2d_uniform(d.nx, d.ny, d.dx, d.dy);
    else
      A = an ? lap3d_aniso(d.nx, d.ny, d.nz, d.dx, c.data()) : lap3d_uniform(d.nx, d.ny, d.nz, d.dx);
    built = true;
  }

  // NLSESolver::step nlse_cubic_solver.hpp:55-74; NLSEQuinticSolver::step
  // nlse_cubic_quintic_solver.hpp:10-29 (post-phase density from the OLD u, :23-24);
  // NLSESaturatingSolver::step nlse_saturating_solver.hpp:11-33 (complex denominator).
  bool step_nlse(cplx *u, double dt) {
    const cplx tau(0, dt);
    std::vector<cplx> rho(n), buf(n);
    auto g = [&](cplx uu, cplx den_src, int64_t i) -> cplx {
      const double r = uu.real() * uu.real() + uu.imag() * uu.imag();
      switch (d.equation) {
      case NLS_EQ_NLSE_CUBIC: return cplx(m[i] * r);
      case NLS_EQ_NLSE_CQ: return cplx(m[i] * (d.s1 * r + d.s2 * (r * r)));
      default: return m[i] * (cplx(r) * (cplx(1.0) / (cplx(1.0) + d.kappa * den_src)));
      }
    };
    for (int64_t i = 0; i < n; ++i) rho[i] = std::exp(.5 * tau * g(u[i], u[i], i)) * u[i];
    if (!matfunc_apply(A, sgn, NLS_FN_EXP, tau, rho.data(), (int)d.krylov_m, d.reorth, buf.data()))
      return false;
    for (int64_t i = 0; i < n; ++i) {
      // cubic: density from buf; cq/sat: |u_old|^2 (sat: denominator from buf)
      const cplx src = d.equation == NLS_EQ_NLSE_CUBIC ? buf[i] : u[i];
      u[i] = std::exp(.5 * tau * g(src, buf[i], i)) * buf[i];
    }
    return true;
  }

  // sEWI step.  Default convention (NLSEGautschiSolver3d::step nlse_cubic_gautschi_solver_3d.hpp:8-35,
  // NLSECubicQuinticGautschiSolver::step nlse_cubic_quintic_gautschi_solver.hpp:11-41,
  // NLSESaturatingGautschiSolver::step nlse_saturating_gautschi_solver.hpp:11-44 [real denominator
  // 1 + kappa |u|^2 here], device step_sewi nlse_dev.hpp:205-238):
  //   u+ = expm(L, u-, 2 tau) - 2 tau expm(L, sincm(L, B(u), dt), tau)
  // NLS_FLAG_SEWI_HOST2D (NLSECubicGautschiSolver::step nlse_cubic_gautschi_solver.hpp:21-39):
  //   u+ = expm(L, u-, -2 tau) + 2 tau expm(L, sincm(L, B(u), dt), -tau)
  // First step: one SS2 step with u- = u0 (nlse_cubic_gautschi_driver.cpp:122-125).
  bool step_sewi(cplx *u, cplx *uprev, char &started, double dt) {
    if (!started) {
      std::copy(u, u + n, uprev);
      started = 1;
      return step_nlse(u, dt);
    }
    const cplx tau(0, dt);
    const double sg = (d.flags & NLS_FLAG_SEWI_HOST2D) ? -1.0 : 1.0;
    const int mk = (int)d.krylov_m;
    std::vector<cplx> B(n), phi(n), a(n), b(n);
    for (int64_t i = 0; i < n; ++i) {
      const double r = u[i].real() * u[i].real() + u[i].imag() * u[i].imag();
      double gg;
      switch (d.equation) {
      case NLS_EQ_NLSE_CUBIC: gg = m[i] * r; break;
      case NLS_EQ_NLSE_CQ: gg = m[i] * (d.s1 * r + d.s2 * (r * r)); break;
      default: gg = m[i] * (r * (1.0 / (1.0 + d.kappa * r))); break;
      }
      B[i] = -gg * u[i];
    }
    if (!matfunc_apply(A, sgn, NLS_FN_SINC, cplx(dt, 0.0), B.data(), mk, d.reorth, phi.data())) return false;
    if (!matfunc_apply(A, sgn, NLS_FN_EXP, sg * tau, phi.data(), mk, d.reorth, a.data())) return false;
    if (!matfunc_apply(A, sgn, NLS_FN_EXP, sg * 2. * tau, uprev, mk, d.reorth, b.data())) return false;
    for (int64_t i = 0; i < n; ++i) {
      const cplx un = b[i] - sg * 2. * tau * a[i];
      uprev[i] = u[i];
      u[i] = un;
    }
    return true;
  }

  // SGESolver::step sg_single_solver.hpp:43-59 (filter mod_cosine) and the id_sqrt-filter
  // family kg_solver.hpp:13-22, phi4_solver.hpp, sg_double_solver.hpp, sg_hyperbolic_solver.hpp
  bool step_gautschi(double *u, double *up, double tau) {
    const int mk = (int)d.krylov_m;
    std::vector<double> b2(n), b3(n), cs(n);
    const int filt = d.equation == NLS_EQ_SGE ? NLS_FN_MOD_COSINE : NLS_FN_ID_SQRT;
    if (!matfunc_apply(A, sgn, filt, cplx(tau), u, mk, d.reorth, b2.data())) return false;
    for (int64_t i = 0; i < n; ++i) b2[i] = -(m[i] * f_real(d.equation, b2[i]));
    if (!matfunc_apply(A, sgn, NLS_FN_SINC2_HALF, cplx(tau), b2.data(), mk, d.reorth, b3.data()))
      return false;
    if (!matfunc_apply(A, sgn, NLS_FN_COS_SQRT, cplx(tau), u, mk, d.reorth, cs.data())) return false;
    for (int64_t i = 0; i < n; ++i) {
      const double un = 2 * cs[i] - up[i] + tau * tau * b3[i];
      up[i] = u[i];
      u[i] = un;
    }
    return true;
  }

  // SGESolverSV::step sg_single_sv_solver.hpp:9-19, KGESVSolver::step kg_sv_solver.hpp:13-21,
  // SGEHyperbolicSVSolver::step sg_hyperbolic_sv_solver.hpp:13-22
  bool step_sv(double *u, double *up, double tau) {
    std::vector<double> buf(n);
    spmv(A, sgn, u, buf.data());
    for (int64_t i = 0; i < n; ++i) {
      const double b = buf[i] - m[i] * f_real(d.equation, u[i]);
      const double un = 2 * u[i] - up[i] + tau * tau * b;
      up[i] = u[i];
      u[i] = un;
    }
    return true;
  }
};

extern "C" {

const char *orc_last_error(void) { return g_err.c_str(); }

int orc_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// launchers such as torchrun export OMP_NUM_THREADS=1; the timed CPU baseline asks for all cores explicitly
void orc_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

int orc_create(const nls_desc *desc, orc_handle *out) {
  if (!desc || !out || desc->struct_size != sizeof(nls_desc)) {
    g_err = "bad descriptor";
    return NLS_ERR_ARG;
  }
  if (desc->nx < 3 || desc->ny < 3 || desc->nz == 2 || desc->nz == 0 || desc->krylov_m < 2 ||
      desc->krylov_m > 64 || desc->batch < 1) {
    g_err = "bad sizes";
    return NLS_ERR_ARG;
  }
  auto *e = new orc_engine;
  e->d = *desc;
  e->n = (int64_t)desc->nx * desc->ny * desc->nz;
  e->cplx_state = desc->equation <= NLS_EQ_NLSE_SAT;
  e->sgn = desc->op_sign == 0.0 ? 1.0 : desc->op_sign;
  e->m.assign(e->n, 0.0); // nlse_cubic_driver.cpp:63: m defaults to ZERO
  if (e->cplx_state) {
    e->uc.assign((size_t)e->n * desc->batch, cplx(0));
    e->uc_prev.assign((size_t)e->n * desc->batch, cplx(0));
    e->sewi_started.assign(desc->batch, 0);
  } else {
    e->ur.assign((size_t)e->n * desc->batch, 0.0);
    e->ur_past.assign((size_t)e->n * desc->batch, 0.0);
  }
  *out = e;
  return NLS_OK;
}

int orc_destroy(orc_handle h) {
  delete h;
  return NLS_OK;
}

int orc_set_field(orc_handle e, int field, const void *src, size_t bytes) {
  const size_t sb = (size_t)e->n * e->d.batch * (e->cplx_state ? 16 : 8);
  switch (field) {
  case NLS_FIELD_U:
    if (bytes != sb) break;
    std::memcpy(e->cplx_state ? (void *)e->uc.data() : (void *)e->ur.data(), src, bytes);
    if (e->cplx_state) std::fill(e->sewi_started.begin(), e->sewi_started.end(), 0);
    return NLS_OK;
  case NLS_FIELD_U_PAST:
    if (bytes != sb) break;
    if (e->cplx_state) {
      if (e->d.integrator != NLS_INT_SEWI) break;
      std::memcpy(e->uc_prev.data(), src, bytes);
      std::fill(e->sewi_started.begin(), e->sewi_started.end(), 1);
      return NLS_OK;
    }
    std::memcpy(e->ur_past.data(), src, bytes);
    return NLS_OK;
  case NLS_FIELD_V: { // u_past = u0 - dt*v0  sg_single_driver.cpp:70
    if (e->cplx_state || bytes != sb) break;
    const double *v = (const double *)src;
    for (size_t i = 0; i < e->ur.size(); ++i) e->ur_past[i] = e->ur[i] - e->d.dt * v[i];
    return NLS_OK;
  }
  case NLS_FIELD_M:
    if (bytes != (size_t)e->n * 8) break;
    std::memcpy(e->m.data(), src, bytes);
    return NLS_OK;
  case NLS_FIELD_C:
    if (bytes != (size_t)e->n * 8) break;
    e->c.resize(e->n);
    std::memcpy(e->c.data(), src, bytes);
    e->built = false;
    return NLS_OK;
  }
  g_err = "set_field: bad field or size";
  return NLS_ERR_ARG;
}

int orc_get_field(orc_handle e, int field, void *dst, size_t bytes) {
  const size_t sb = (size_t)e->n * e->d.batch * (e->cplx_state ? 16 : 8);
  switch (field) {
  case NLS_FIELD_U:
    if (bytes != sb) break;
    std::memcpy(dst, e->cplx_state ? (void *)e->uc.data() : (void *)e->ur.data(), bytes);
    return NLS_OK;
  case NLS_FIELD_U_PAST:
    if (bytes != sb) break;
    if (e->cplx_state) {
      if (e->d.integrator != NLS_INT_SEWI) break;
      std::memcpy(dst, e->uc_prev.data(), bytes);
      return NLS_OK;
    }
    std::memcpy(dst, e->ur_past.data(), bytes);
    return NLS_OK;
  case NLS_FIELD_V: { // v = (u - u_past)/dt  sg_single_driver.cpp:113
    if (e->cplx_state || bytes != sb) break;
    double *v = (double *)dst;
    for (size_t i = 0; i < e->ur.size(); ++i) v[i] = (e->ur[i] - e->ur_past[i]) / e->d.dt;
    return NLS_OK;
  }
  }
  g_err = "get_field: bad field or size";
  return NLS_ERR_ARG;
}

// driver loop body: step + neumann_bc_no_velocity (nlse_cubic_driver.cpp:96-98,
// sg_single_driver.cpp:110-112).  Batch = independent trajectories (one per OpenMP thread,
// the analogue of the reference's one-process-per-trajectory launch, submit_nlse.py:117-129).
int orc_step(orc_handle e, double dt, uint32_t nsteps) {
  e->build();
  int bad = 0;
  const int64_t n = e->n;
#pragma omp parallel for schedule(dynamic) reduction(+ : bad)
  for (int b = 0; b < (int)e->d.batch; ++b) {
    for (uint32_t s = 0; s < nsteps; ++s) {
      bool ok;
      if (e->cplx_state) {
        cplx *u = &e->uc[(size_t)b * n];
        if (e->d.integrator == NLS_INT_SEWI)
          ok = e->step_sewi(u, &e->uc_prev[(size_t)b * n], e->sewi_started[b], dt);
        else ok = e->step_nlse(u, dt);
        neumann_bc(u, e->d.nx, e->d.ny, e->d.nz);
      } else {
        double *u = &e->ur[(size_t)b * n], *up = &e->ur_past[(size_t)b * n];
        ok = e->d.integrator == NLS_INT_SV ? e->step_sv(u, up, dt) : e->step_gautschi(u, up, dt);
        neumann_bc(u, e->d.nx, e->d.ny, e->d.nz);
      }
      if (!ok) {
        bad++;
        break;
      }
    }
  }
  if (bad) {
    g_err = "tridiagonal eigensolver did not converge";
    return NLS_ERR_STATE;
  }
  return NLS_OK;
}

int orc_stencil_apply(orc_handle e, const void *x, void *y) {
  e->build();
  for (uint32_t b = 0; b < e->d.batch; ++b) {
    if (e->cplx_state) spmv(e->A, e->sgn, (const cplx *)x + (size_t)b * e->n, (cplx *)y + (size_t)b * e->n);
    else spmv(e->A, e->sgn, (const double *)x + (size_t)b * e->n, (double *)y + (size_t)b * e->n);
  }
  return NLS_OK;
}

int orc_matfunc(orc_handle e, int fn, double t_re, double t_im, const void *in, void *out) {
  e->build();
  bool ok = true;
  for (uint32_t b = 0; b < e->d.batch && ok; ++b) {
    if (e->cplx_state)
      ok = matfunc_apply(e->A, e->sgn, fn, cplx(t_re, t_im), (const cplx *)in + (size_t)b * e->n,
                         (int)e->d.krylov_m, e->d.reorth, (cplx *)out + (size_t)b * e->n);
    else
      ok = matfunc_apply(e->A, e->sgn, fn, cplx(t_re, t_im), (const double *)in + (size_t)b * e->n,
                         (int)e->d.krylov_m, e->d.reorth, (double *)out + (size_t)b * e->n);
  }
  return ok ? NLS_OK : NLS_ERR_STATE;
}

int orc_lanczos(orc_handle e, const void *in, void *V, void *alpha, double *beta) {
  e->build();
  const int m = (int)e->d.krylov_m;
  std::vector<double> off;
  double b0;
  if (e->cplx_state) {
    std::vector<cplx> Vv, a;
    b0 = lanczos(e->A, e->sgn, (const cplx *)in, m, e->d.reorth, Vv, a, off);
    if (V) std::memcpy(V, Vv.data(), Vv.size() * 16);
    if (alpha) std::memcpy(alpha, a.data(), a.size() * 16);
  } else {
    std::vector<double> Vv, a;
    b0 = lanczos(e->A, e->sgn, (const double *)in, m, e->d.reorth, Vv, a, off);
    if (V) std::memcpy(V, Vv.data(), Vv.size() * 8);
    if (alpha) std::memcpy(alpha, a.data(), a.size() * 8);
  }
  if (beta) {
    beta[0] = b0;
    for (int j = 0; j + 1 < m; ++j) beta[j + 1] = off[j];
  }
  return NLS_OK;
}

// numpy.gradient along one axis (first-order one-sided at both ends, centred inside), times 1:
// returns (f[p+] - f[p-]) / (h * #steps)                      energy_sge.py:7-8
static double np_gradient(const double *f, int64_t pos, int64_t len, int64_t stride, double h) {
  const int64_t lo = pos > 0 ? 1 : 0, hi = pos < len - 1 ? 1 : 0;
  return (f[hi * stride] - f[-lo * stride]) / (h * (double)(lo + hi));
}

int orc_reduce(orc_handle e, int what, double *out) {
  const double cell = e->d.nz == 1 ? e->d.dx * e->d.dy : e->d.dx * e->d.dy * e->d.dz;
  const int64_t nx = e->d.nx, ny = e->d.ny, nz = e->d.nz, pl = nx * ny;
  if (what == NLS_RED_ENERGY) {
    if (!e->cplx_state && !(e->d.dt > 0)) return NLS_ERR_STATE;
    for (uint32_t b = 0; b < e->d.batch; ++b) {
      double s = 0.0;
      if (e->cplx_state) {
        // finalized_scripts/compare_utils_complex_2d.py:134-151: centred differences on the interior
        // sites only (ux: x neighbours / 2dx, uy: y neighbours / 2dy), potential -1/2 |u|^4, m_eff = 1
        const cplx *u = &e->uc[(size_t)b * e->n];
        for (int64_t k = 0; k < nz; ++k)
          for (int64_t j = 0; j < ny; ++j)
            for (int64_t i = 0; i < nx; ++i) {
              const int64_t p = (k * ny + j) * nx + i;
              const double a2 = abs2_s(u[p]);
              s += -0.5 * a2 * a2;
              if (i == 0 || i == nx - 1 || j == 0 || j == ny - 1 || (nz > 1 && (k == 0 || k == nz - 1))) continue;
              s += abs2_s(u[p + 1] - u[p - 1]) / (4.0 * e->d.dx * e->d.dx);
              s += abs2_s(u[p + nx] - u[p - nx]) / (4.0 * e->d.dy * e->d.dy);
              if (nz > 1) s += abs2_s(u[p + pl] - u[p - pl]) / (4.0 * e->d.dz * e->d.dz);
            }
        out[b] = s * cell;
      } else {
        // energy_sge.py:6-14: np.gradient(u, dx, axis=0), np.gradient(u, dy, axis=1) on the (ny,nx)
        // array -- dx goes with the y index and dy with the x index (kept); 3D: (dz,dy,dx) for (z,y,x)
        const double *u = &e->ur[(size_t)b * e->n], *up = &e->ur_past[(size_t)b * e->n];
        const double hx = nz == 1 ? e->d.dy : e->d.dx, hy = nz == 1 ? e->d.dx : e->d.dy;
        for (int64_t k = 0; k < nz; ++k)
          for (int64_t j = 0; j < ny; ++j)
            for (int64_t i = 0; i < nx; ++i) {
              const int64_t p = (k * ny + j) * nx + i;
              const double gx = np_gradient(u + p, i, nx, 1, hx), gy = np_gradient(u + p, j, ny, nx, hy);
              double g2 = gx * gx + gy * gy;
              if (nz > 1) { const double gz = np_gradient(u + p, k, nz, pl, e->d.dz); g2 += gz * gz; }
              const double v = (u[p] - up[p]) / e->d.dt; // sg_single_driver.cpp:113
              s += 0.5 * (g2 + v * v + 2.0 * (1.0 - std::cos(u[p])));
            }
        out[b] = 0.5 * cell * s;
      }
    }
    return NLS_OK;
  }
  if (what != NLS_RED_MASS && what != NLS_RED_NONFINITE) return NLS_ERR_ARG;
  for (uint32_t b = 0; b < e->d.batch; ++b) {
    double s = 0.0;
    for (int64_t i = 0; i < e->n; ++i) {
      const size_t k = (size_t)b * e->n + i;
      if (what == NLS_RED_MASS) s += e->cplx_state ? abs2_s(e->uc[k]) : e->ur[k] * e->ur[k];
      else {
        const bool fin = e->cplx_state ? (std::isfinite(e->uc[k].real()) && std::isfinite(e->uc[k].imag()))
                                       : std::isfinite(e->ur[k]);
        s += fin ? 0.0 : 1.0;
      }
    }
    out[b] = what == NLS_RED_MASS ? s * cell : s;
  }
  return NLS_OK;
}

} // extern "C"

// finalized_scripts/downsampling.py:68-104 (2D) and :215-258 (3D): RegularGridInterpolator(method="linear")
// from linspace(-L,L,n) onto linspace(-L,L,t) along every axis = multilinear interpolation at index position
// k*(n-1)/(t-1).  dst: [batch][tnz][tny][tnx].  field: NLS_FIELD_U, or NLS_FIELD_V (real states).
template <class T, class F>
static void resample_state(const orc_engine *e, F at, uint32_t tx, uint32_t ty, uint32_t tz, T *dst) {
  const int64_t nx = e->d.nx, ny = e->d.ny, nz = e->d.nz;
  auto axis = [](int64_t a, int64_t t, int64_t n, int64_t &i0, double &f) {
    const double pos = t > 1 ? (double)a * (double)(n - 1) / (double)(t - 1) : 0.0;
    i0 = std::max<int64_t>(0, std::min<int64_t>((int64_t)pos, n - 2));
    f = n > 1 ? pos - (double)i0 : 0.0;
  };
  for (uint32_t b = 0; b < e->d.batch; ++b)
    for (int64_t c = 0; c < tz; ++c)
      for (int64_t bb = 0; bb < ty; ++bb)
        for (int64_t a = 0; a < tx; ++a) {
          int64_t i, j, k;
          double fx, fy, fz;
          axis(a, tx, nx, i, fx); axis(bb, ty, ny, j, fy); axis(c, tz, nz, k, fz);
          const int64_t dxs = nx > 1 ? 1 : 0, dys = ny > 1 ? nx : 0, dzs = nz > 1 ? nx * ny : 0;
          const int64_t p = (int64_t)b * e->n + (k * ny + j) * nx + i;
          auto plane = [&](int64_t q) {
            const T r0 = at(q) * (1.0 - fx) + at(q + dxs) * fx;
            const T r1 = at(q + dys) * (1.0 - fx) + at(q + dys + dxs) * fx;
            return r0 * (1.0 - fy) + r1 * fy;
          };
          T v = plane(p);
          if (nz > 1) v = v * (1.0 - fz) + plane(p + dzs) * fz;
          dst[(((size_t)b * tz + c) * ty + bb) * tx + a] = v;
        }
}

extern "C" {

int orc_get_field_resampled(orc_handle e, int field, uint32_t tx, uint32_t ty, uint32_t tz, void *dst, size_t bytes) {
  if (field != NLS_FIELD_U && field != NLS_FIELD_V) return NLS_ERR_ARG;
  if (tx < 2 || ty < 2 || (e->d.nz == 1 ? tz != 1 : tz < 2)) return NLS_ERR_ARG;
  if (bytes != (size_t)tx * ty * tz * e->d.batch * (e->cplx_state ? 16 : 8)) return NLS_ERR_ARG;
  if (e->cplx_state) {
    if (field != NLS_FIELD_U) return NLS_ERR_STATE;
    resample_state<cplx>(e, [&](int64_t q) { return e->uc[q]; }, tx, ty, tz, (cplx *)dst);
  } else if (field == NLS_FIELD_U) {
    resample_state<double>(e, [&](int64_t q) { return e->ur[q]; }, tx, ty, tz, (double *)dst);
  } else {
    if (!(e->d.dt > 0)) return NLS_ERR_STATE;
    const double inv_dt = 1.0 / e->d.dt;
    resample_state<double>(e, [&](int64_t q) { return (e->ur[q] - e->ur_past[q]) * inv_dt; }, tx, ty, tz, (double *)dst);
  }
  return NLS_OK;
}

int orc_csr(orc_handle e, int64_t *nnz, int64_t *rowptr, int32_t *col, double *val) {
  e->build();
  if (nnz) *nnz = (int64_t)e->A.val.size();
  if (rowptr) std::memcpy(rowptr, e->A.rowptr.data(), e->A.rowptr.size() * 8);
  if (col) std::memcpy(col, e->A.col.data(), e->A.col.size() * 4);
  if (val)
    for (size_t k = 0; k < e->A.val.size(); ++k) val[k] = e->sgn * e->A.val[k];
  return NLS_OK;
}

} // extern "C"
