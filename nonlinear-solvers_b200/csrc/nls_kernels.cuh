// nls_kernels.cuh -- the hot-path kernels (all fp64 SIMT, HBM/L2-bandwidth bound).
//
//   K1  k_stencil            y = L x                         (replaces device/spmv.hpp:65-72)
//   K2  k_lanczos_pa/pb/pc   one Lanczos iteration in three fused passes, scalars stay on
//                            the device (replaces device/lanczos_complex.hpp:318-551 and the
//                            host algorithm eigen_krylov_complex.hpp:13-52 / _real.hpp:8-55)
//   K3  k_tridiag            m x m tridiagonal eigensolve + f(Lambda) -> coefficient vectors
//                            (replaces cusolverDn{D,Z}{sy,he}evd, device/matfunc_*.hpp)
//   K4  k_nlse_finish, k_gautschi_a/b, k_combine
//                            basis recombination fused with the pointwise nonlinear step, the
//                            no-flux boundary copy, the next start vector and its norm
//   K5  k_sv                 Stoermer-Verlet: stencil + update in one kernel
//
// Per-trajectory device scalars ("scalar block", doubles):
//   [0 .. 64]           nrm2[j]  : nrm2[0] = ||start||^2, nrm2[j+1] = ||w_j||^2 = T(j+1,j)^2
//   [65 .. 65+128)      alpha[j] : T(j,j) (W doubles each)
//   [193 .. 193+128)    h[i]     : re-orthogonalisation coefficients of the current iteration
#pragma once
#include "nls_device.cuh"
#include "../../include/nls.h"

namespace nls {

constexpr int kOffNrm2 = 0;
constexpr int kOffAlpha = 65;
constexpr int kOffH = 193;
constexpr int kScalStride = 328;
constexpr int kCoefStride = 2 * kMaxM; // doubles per coefficient vector

// arrays hold [batch][stride] scalars with `halo` ghost elements in front of site 0
template <class S>
__device__ __forceinline__ S *traj(S *base, const Geom &g) {
  return base + (size_t)blockIdx.y * g.stride + g.halo;
}
template <class S>
__device__ __forceinline__ const S *traj(const S *base, const Geom &g) {
  return base + (size_t)blockIdx.y * g.stride + g.halo;
}

// ------------------------------------------------------------------ K1
template <class S, bool ANISO>
__global__ void __launch_bounds__(kThreads)
k_stencil(Geom g, const S *__restrict__ x, S *__restrict__ y, const double *__restrict__ cf) {
  const S *xb = traj(x, g);
  S *yb = traj(y, g);
  const double *cb = ANISO ? cf + g.halo : nullptr;
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads)
    yb[idx] = stencil_site<S, ANISO>(g, xb, cb, idx, 1.0);
}

// ||x||^2 -> dst[b*dst_stride]
template <class S>
__global__ void __launch_bounds__(kThreads)
k_norm2(Geom g, const S *__restrict__ x, double *dst, int dst_stride, RedCtx rc) {
  const S *xb = traj(x, g);
  double acc[1] = {0.0};
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads)
    acc[0] += Sc<S>::abs2(xb[idx]);
  grid_reduce<1>(acc, rc, dst + (size_t)blockIdx.y * dst_stride);
}

// ------------------------------------------------------------------ K2, pass a
// v_j = wprev / beta ; w = L v_j - beta v_{j-1} ; alpha_j = v_j^H w
// (eigen_krylov_complex.hpp:22,:25-28,:46; the normalisation of the previous iteration is
//  fused in so the un-normalised vector is read once, with its stencil halo).
template <class S, bool ANISO>
__global__ void __launch_bounds__(kThreads)
k_lanczos_pa(Geom g, const S *__restrict__ wprev, const S *__restrict__ vprev, S *__restrict__ vj,
             S *__restrict__ wout, const double *__restrict__ cf, double *scal, int j, RedCtx rc) {
  using T = Sc<S>;
  double *sb = scal + (size_t)blockIdx.y * kScalStride;
  const double beta = sqrt(sb[kOffNrm2 + j]);
  const double inv = 1.0 / beta;
  const S *wp = traj(wprev, g);
  const S *vp = j > 0 ? traj(vprev, g) : nullptr;
  S *vo = traj(vj, g);
  S *wo = traj(wout, g);
  const double *cb = ANISO ? cf + g.halo : nullptr;
  double acc[T::W];
#pragma unroll
  for (int k = 0; k < T::W; ++k) acc[k] = 0.0;
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    const S vc = T::mulr(wp[idx], inv);
    S w = stencil_site<S, ANISO>(g, wp, cb, idx, inv);
    if (j > 0) w = T::fma_r(-beta, vp[idx], w);
    vo[idx] = vc;
    wo[idx] = w;
    T::dot_acc(acc, vc, w);
  }
  grid_reduce<T::W>(acc, rc, sb + kOffAlpha + j * T::W);
}

// ------------------------------------------------------------------ K2, pass b
// w -= alpha_j v_j (first chunk only) ; h_i = v_i^H w for i in [i0, i0+cnt)
// (eigen_krylov_complex.hpp:29 and the dot products of the sweep :32-33, batched)
template <class S, int NV>
__global__ void __launch_bounds__(kThreads)
k_lanczos_pb(Geom g, S *__restrict__ w, const S *__restrict__ V, size_t colstride, int j, int i0,
             int cnt, double *scal, RedCtx rc) {
  using T = Sc<S>;
  double *sb = scal + (size_t)blockIdx.y * kScalStride;
  const S alpha = T::from(sb + kOffAlpha + j * T::W);
  S *wb = traj(w, g);
  const S *vb = traj(V, g);
  double acc[NV * T::W];
#pragma unroll
  for (int k = 0; k < NV * T::W; ++k) acc[k] = 0.0;
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    S wv = wb[idx];
    if (i0 == 0) {
      wv = T::sub(wv, T::mul(alpha, vb[(size_t)j * colstride + idx]));
      wb[idx] = wv;
    }
#pragma unroll
    for (int i = 0; i < NV; ++i)
      if (i < cnt) T::dot_acc(acc + i * T::W, vb[(size_t)(i0 + i) * colstride + idx], wv);
  }
  grid_reduce<NV * T::W>(acc, rc, sb + kOffH + i0 * T::W);
}

// ------------------------------------------------------------------ K2, pass c
// w -= sum_i h_i v_i for i in [i0, i0+cnt) ; if LAST: ||w||^2 -> nrm2[j+1]
// (eigen_krylov_complex.hpp:34 batched, :38)
template <class S, int NV, bool LAST>
__global__ void __launch_bounds__(kThreads)
k_lanczos_pc(Geom g, S *__restrict__ w, const S *__restrict__ V, size_t colstride, int j, int i0,
             int cnt, const double *hsrc_base, int hoff, double *scal, RedCtx rc, Halo<S> hp) {
  using T = Sc<S>;
  double *sb = scal + (size_t)blockIdx.y * kScalStride;
  const double *hs = hsrc_base + (size_t)blockIdx.y * kScalStride + hoff;
  S h[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) h[i] = i < cnt ? T::from(hs + i * T::W) : T::zero();
  S *wb = traj(w, g);
  const S *vb = traj(V, g);
  double acc[1] = {0.0};
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    S wv = wb[idx];
#pragma unroll
    for (int i = 0; i < NV; ++i)
      if (i < cnt) wv = T::sub(wv, T::mul(h[i], vb[(size_t)(i0 + i) * colstride + idx]));
    wb[idx] = wv;
    if (LAST) {
      acc[0] += T::abs2(wv);
      halo_push<S>(hp, g.n, g.plane, idx, wv); // w_j is the next iteration's stencil input
    }
  }
  if (LAST) grid_reduce<1>(acc, rc, sb + kOffNrm2 + j + 1);
}

// ------------------------------------------------------------------ K3
// One warp per trajectory: implicit-shift QL on the m x m symmetric tridiagonal T
// (diag = Re alpha_j, alpha_{m-1} = 0 as in the reference; off-diag = sqrt(nrm2[j+1])),
// eigenvector rows distributed over the lanes, then
//   coef_i = beta0 * sum_k Z_ik f(t, lambda_k) Z_0k            (beta * V * f(T) * e1, :77-82)
// with the last coefficient divided by T(m-1,m-2) because column m-1 of the basis is kept
// un-normalised in the w buffer.  Function tables: eigen_krylov_complex.hpp:70-76,:89-96,
// eigen_krylov_real.hpp:79-80,:109-115,:142-144,:161-170,:189-204.
__device__ __forceinline__ double2 fn_eval(int fn, double tr, double ti, double lam) {
  switch (fn) {
  case NLS_FN_EXP: { // exp((tr + i ti) lam)
    double s, c;
    sincos(ti * lam, &s, &c);
    const double e = exp(tr * lam);
    return make_double2(e * c, e * s);
  }
  case NLS_FN_SINC: { // real-valued t in every call site (sEWI): x = tr*lam
    const double x = tr * lam;
    return make_double2(fabs(x) < 1e-8 ? 1.0 : sin(x) / x, 0.0);
  }
  case NLS_FN_COS_SQRT: return make_double2(cos(tr * sqrt(fabs(lam))), 0.0);
  case NLS_FN_SINC2_SQRT: {
    const double x = tr * sqrt(fabs(lam));
    const double s = fabs(x) < 1e-8 ? 1.0 : sin(x) / x;
    return make_double2(s * s, 0.0);
  }
  case NLS_FN_ID_SQRT: return make_double2(tr * sqrt(fabs(lam)), 0.0);
  case NLS_FN_SINC2_HALF: {
    const double x = tr / 2. * sqrt(fabs(lam));
    const double s = fabs(x) < 1e-8 ? 1.0 : sin(x) / x;
    return make_double2(s * s, 0.0);
  }
  default: { // NLS_FN_MOD_COSINE
    const double th = tr * sqrt(fabs(lam));
    double v = 1.0;
    if (fabs(th) >= 1e-12) {
      const double ch = cos(th / 2.0);
      v = ch * ch * sin(th) / th;
    }
    return make_double2(v * v, 0.0);
  }
  }
}

struct TridiagArgs {
  int m, W, nfn;
  int fn[2];
  double tr[2], ti[2];
};

__global__ void __launch_bounds__(32)
k_tridiag(const double *__restrict__ scal, double *__restrict__ coef, TridiagArgs a, int *status) {
  __shared__ double d[kMaxM], e[kMaxM], Z[kMaxM][kMaxM + 1];
  const int lane = threadIdx.x, m = a.m;
  const double *sb = scal + (size_t)blockIdx.x * kScalStride;
  for (int i = lane; i < m; i += 32) {
    d[i] = i < m - 1 ? sb[kOffAlpha + i * a.W] : 0.0;
    e[i] = i < m - 1 ? sqrt(sb[kOffNrm2 + i + 1]) : 0.0;
    for (int k = 0; k < m; ++k) Z[i][k] = i == k ? 1.0 : 0.0;
  }
  __syncwarp();
  bool ok = true;
  for (int l = 0; l < m && ok; ++l) {
    int iter = 0, mm;
    do {
      for (mm = l; mm < m - 1; ++mm) {
        const double dd = fabs(d[mm]) + fabs(d[mm + 1]);
        if (fabs(e[mm]) <= 2.220446049250313e-16 * dd) break;
      }
      if (mm != l) {
        if (iter++ == 60) { ok = false; break; }
        double g = (d[l + 1] - d[l]) / (2.0 * e[l]);
        double r = hypot(g, 1.0);
        g = d[mm] - d[l] + e[l] / (g + copysign(r, g));
        double s = 1.0, c = 1.0, p = 0.0;
        int i;
        bool early = false;
        for (i = mm - 1; i >= l; --i) {
          double f = s * e[i];
          const double b = c * e[i];
          r = hypot(f, g);
          const double di = d[i], di1 = d[i + 1];
          __syncwarp();
          if (lane == 0) e[i + 1] = r;
          if (r == 0.0) {
            if (lane == 0) { d[i + 1] = di1 - p; e[mm] = 0.0; }
            early = true;
            break;
          }
          s = f / r;
          c = g / r;
          g = di1 - p;
          r = (di - g) * s + 2.0 * c * b;
          p = s * r;
          if (lane == 0) d[i + 1] = g + p;
          g = c * r - b;
          for (int k = lane; k < m; k += 32) { // rotate columns i, i+1 of Z (rows over lanes)
            f = Z[k][i + 1];
            Z[k][i + 1] = s * Z[k][i] + c * f;
            Z[k][i] = c * Z[k][i] - s * f;
          }
        }
        __syncwarp();
        if (early) continue;
        if (lane == 0) { d[l] -= p; e[l] = g; e[mm] = 0.0; }
        __syncwarp();
      }
    } while (mm != l);
  }
  __syncwarp();
  if (!ok && lane == 0) atomicExch(status, 1);
  const double beta0 = sqrt(sb[kOffNrm2 + 0]);
  const double last = 1.0 / sqrt(sb[kOffNrm2 + m - 1]); // 1 / T(m-1,m-2)
  __shared__ double2 fv[kMaxM];
  for (int f = 0; f < a.nfn; ++f) {
    __syncwarp();
    for (int k = lane; k < m; k += 32) fv[k] = fn_eval(a.fn[f], a.tr[f], a.ti[f], d[k]);
    __syncwarp();
    double *co = coef + ((size_t)blockIdx.x * 2 + f) * kCoefStride;
    for (int i = lane; i < m; i += 32) {
      double sr = 0.0, si = 0.0;
      for (int k = 0; k < m; ++k) {
        const double zz = Z[i][k] * Z[0][k];
        sr = fma(zz, fv[k].x, sr);
        si = fma(zz, fv[k].y, si);
      }
      double scale = beta0;
      if (i == m - 1) scale *= last;
      co[2 * i] = sr * scale;
      co[2 * i + 1] = si * scale;
    }
  }
}

// ------------------------------------------------------------------ K4 helpers
// y(idx) = sum_{i<m-1} coef_i V_i[idx] + coef_{m-1} wlast[idx]
template <class S>
__device__ __forceinline__ S combine_site(const S *__restrict__ vb, size_t colstride,
                                          const S *__restrict__ wl, const double *__restrict__ co,
                                          int m, long long idx);
template <>
__device__ __forceinline__ double combine_site<double>(const double *__restrict__ vb, size_t colstride,
                                                       const double *__restrict__ wl,
                                                       const double *__restrict__ co, int m,
                                                       long long idx) {
  double s = 0.0;
  for (int i = 0; i < m - 1; ++i) s = fma(vb[(size_t)i * colstride + idx], co[2 * i], s);
  return fma(wl[idx], co[2 * (m - 1)], s);
}
template <>
__device__ __forceinline__ double2 combine_site<double2>(const double2 *__restrict__ vb,
                                                         size_t colstride,
                                                         const double2 *__restrict__ wl,
                                                         const double *__restrict__ co, int m,
                                                         long long idx) {
  double2 s = make_double2(0.0, 0.0);
  for (int i = 0; i < m - 1; ++i) {
    const double2 v = vb[(size_t)i * colstride + idx];
    const double cr = co[2 * i], ci = co[2 * i + 1];
    s.x = fma(v.x, cr, fma(-v.y, ci, s.x));
    s.y = fma(v.x, ci, fma(v.y, cr, s.y));
  }
  const double2 v = wl[idx];
  const double cr = co[2 * (m - 1)], ci = co[2 * (m - 1) + 1];
  s.x = fma(v.x, cr, fma(-v.y, ci, s.x));
  s.y = fma(v.x, ci, fma(v.y, cr, s.y));
  return s;
}

// out = sum coef_i v_i   (nls_matfunc: beta * V * f(T) * e1)
template <class S>
__global__ void __launch_bounds__(kThreads)
k_combine(Geom g, const S *__restrict__ V, size_t colstride, const S *__restrict__ wlast,
          const double *__restrict__ coef, int fsel, int m, S *__restrict__ out) {
  __shared__ double co[kCoefStride];
  const double *cg = coef + ((size_t)blockIdx.y * 2 + fsel) * kCoefStride;
  for (int i = threadIdx.x; i < 2 * m; i += kThreads) co[i] = cg[i];
  __syncthreads();
  const S *vb = traj(V, g), *wl = traj(wlast, g);
  S *ob = traj(out, g);
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads)
    ob[idx] = combine_site<S>(vb, colstride, wl, co, m, idx);
}

// nonlinear half-step phase factor exp(.5*tau*g), tau = i*dt, applied to z.
//   cubic:  g = m |a|^2                    nlse_cubic_solver.hpp:62-66
//   cq:     g = m (s1 |a|^2 + s2 |a|^4)    nlse_cubic_quintic_solver.hpp:16-20
//   sat:    g = m |a|^2 / (1 + kappa*q)    nlse_saturating_solver.hpp:15-22 (q complex: as written)
// a = density source, q = denominator source.
struct NlseParams {
  double hdt; // 0.5*dt
  double s1, s2, kappa;
};
template <int EQ>
__device__ __forceinline__ double2 nlse_phase(const NlseParams &p, double mval, double2 a, double2 q,
                                              double2 z) {
  const double r = a.x * a.x + a.y * a.y;
  double s, c, amp = 1.0;
  if (EQ == NLS_EQ_NLSE_CUBIC) {
    sincos(p.hdt * (mval * r), &s, &c);
  } else if (EQ == NLS_EQ_NLSE_CQ) {
    sincos(p.hdt * (mval * (p.s1 * r + p.s2 * (r * r))), &s, &c);
  } else {
    const double qr = 1.0 + p.kappa * q.x, qi = p.kappa * q.y;
    const double den = qr * qr + qi * qi;
    const double xr = mval * (r * (qr / den)), xi = mval * (r * (-qi / den));
    sincos(p.hdt * xr, &s, &c); // exp((0,hdt)*(xr,xi)) = e^{-hdt*xi} (cos(hdt*xr), sin(hdt*xr))
    amp = exp(-p.hdt * xi);
  }
  c *= amp;
  s *= amp;
  return make_double2(z.x * c - z.y * s, z.x * s + z.y * c);
}

// first half-step of the very first step: rho = exp(.5 tau g(u)) u, ||rho||^2 -> nrm2[0]
template <int EQ>
__global__ void __launch_bounds__(kThreads)
k_nlse_prephase(Geom g, const double2 *__restrict__ u, const double *__restrict__ mf,
                double2 *__restrict__ rho, NlseParams p, double *scal, RedCtx rc, Halo<double2> hp) {
  const double2 *ub = traj(u, g);
  double2 *rb = traj(rho, g);
  const double *mb = mf + g.halo;
  double acc[1] = {0.0};
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    const double2 uv = ub[idx];
    const double2 r = nlse_phase<EQ>(p, mb[idx], uv, uv, uv);
    rb[idx] = r;
    halo_push<double2>(hp, g.n, g.plane, idx, r);
    acc[0] += r.x * r.x + r.y * r.y;
  }
  grid_reduce<1>(acc, rc, scal + (size_t)blockIdx.y * kScalStride + kOffNrm2);
}

// Strang step tail: buf = beta V f(T) e1 ; u = exp(.5 tau g) buf ; no-flux copy ;
// next step's rho = exp(.5 tau g(u)) u and ||rho||^2.   (nlse_cubic_solver.hpp:68-73 +
// boundaries.hpp:40-57 + the next call's :62-66.)  u_old is read at the clamped site, so the
// new state goes to a second buffer (uout) to stay race-free.
template <int EQ>
__global__ void __launch_bounds__(kThreads)
k_nlse_finish(Geom g, const double2 *__restrict__ V, size_t colstride,
              const double2 *__restrict__ wlast, const double *__restrict__ coef, int m,
              const double2 *__restrict__ uold, const double *__restrict__ mf,
              double2 *__restrict__ uout, double2 *__restrict__ rho, NlseParams p, double *scal,
              RedCtx rc, Halo<double2> hp) {
  __shared__ double co[kCoefStride];
  const double *cg = coef + ((size_t)blockIdx.y * 2) * kCoefStride;
  for (int i = threadIdx.x; i < 2 * m; i += kThreads) co[i] = cg[i];
  __syncthreads();
  const double2 *vb = traj(V, g), *wl = traj(wlast, g), *uo = traj(uold, g);
  double2 *un = traj(uout, g), *rb = traj(rho, g);
  const double *mb = mf + g.halo;
  double acc[1] = {0.0};
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    const long long sc = clamp_site(g, idx);
    const double2 buf = combine_site<double2>(vb, colstride, wl, co, m, sc);
    double2 unew;
    if (EQ == NLS_EQ_NLSE_CUBIC) unew = nlse_phase<EQ>(p, mb[sc], buf, buf, buf);
    else unew = nlse_phase<EQ>(p, mb[sc], uo[sc], buf, buf); // density from the OLD u (Q7)
    un[idx] = unew;
    const double2 r = nlse_phase<EQ>(p, mb[idx], unew, unew, unew);
    rb[idx] = r;
    halo_push<double2>(hp, g.n, g.plane, idx, r);
    acc[0] += r.x * r.x + r.y * r.y;
  }
  grid_reduce<1>(acc, rc, scal + (size_t)blockIdx.y * kScalStride + kOffNrm2);
}

// f(u) of the real equations
template <int EQ>
__device__ __forceinline__ double f_real(double x) {
  if (EQ == NLS_EQ_SGE) return sin(x);
  if (EQ == NLS_EQ_KGE) return x * x * x;
  if (EQ == NLS_EQ_PHI4) return (x - x * x * x);
  if (EQ == NLS_EQ_SGE_DOUBLE) return sin(x) + .6 * sin(2. * x);
  return sinh(x);
}

// Gautschi stage A (after the Lanczos run on u): two coefficient sets in one pass over V:
//   filt = beta V phi(T) e1, cosu = beta V cos(T) e1
//   g = -m f(filt)  -> start vector of the second Lanczos run (+ its norm)
//   acc = 2 cosu - u_past ; u_past = u            (sg_single_solver.hpp:52-54,:56-58)
template <int EQ>
__global__ void __launch_bounds__(kThreads)
k_gautschi_a(Geom g, const double *__restrict__ V, size_t colstride,
             const double *__restrict__ wlast, const double *__restrict__ coef, int m,
             const double *__restrict__ u, double *__restrict__ upast, const double *__restrict__ mf,
             double *__restrict__ gout, double *__restrict__ accout, double *scalB, RedCtx rc,
             Halo<double> hp) {
  __shared__ double cof[kCoefStride], coc[kCoefStride];
  const double *cg = coef + ((size_t)blockIdx.y * 2) * kCoefStride;
  for (int i = threadIdx.x; i < 2 * m; i += kThreads) {
    cof[i] = cg[i];
    coc[i] = cg[kCoefStride + i];
  }
  __syncthreads();
  const double *vb = traj(V, g), *wl = traj(wlast, g), *ub = traj(u, g);
  double *up = traj(upast, g), *go = traj(gout, g), *ao = traj(accout, g);
  const double *mb = mf + g.halo;
  double acc[1] = {0.0};
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    double sf = 0.0, sc = 0.0;
    for (int i = 0; i < m - 1; ++i) {
      const double v = vb[(size_t)i * colstride + idx];
      sf = fma(v, cof[2 * i], sf);
      sc = fma(v, coc[2 * i], sc);
    }
    const double v = wl[idx];
    sf = fma(v, cof[2 * (m - 1)], sf);
    sc = fma(v, coc[2 * (m - 1)], sc);
    const double gv = -(mb[idx] * f_real<EQ>(sf));
    go[idx] = gv;
    halo_push<double>(hp, g.n, g.plane, idx, gv);
    acc[0] += gv * gv;
    ao[idx] = 2 * sc - up[idx];
    up[idx] = ub[idx];
  }
  grid_reduce<1>(acc, rc, scalB + (size_t)blockIdx.y * kScalStride + kOffNrm2);
}

// Gautschi stage B: u = acc + tau^2 * beta V sinc2(T) e1, no-flux copy, ||u||^2 for the next step
template <int DUMMY>
__global__ void __launch_bounds__(kThreads)
k_gautschi_b(Geom g, const double *__restrict__ V, size_t colstride,
             const double *__restrict__ wlast, const double *__restrict__ coef, int m,
             const double *__restrict__ accin, double *__restrict__ u, double tau2, double *scalA,
             RedCtx rc, Halo<double> hp) {
  __shared__ double co[kCoefStride];
  const double *cg = coef + ((size_t)blockIdx.y * 2) * kCoefStride;
  for (int i = threadIdx.x; i < 2 * m; i += kThreads) co[i] = cg[i];
  __syncthreads();
  const double *vb = traj(V, g), *wl = traj(wlast, g), *ai = traj(accin, g);
  double *ub = traj(u, g);
  double acc[1] = {0.0};
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    const long long sc = clamp_site(g, idx);
    const double s = combine_site<double>(vb, colstride, wl, co, m, sc);
    const double un = ai[sc] + tau2 * s;
    ub[idx] = un;
    halo_push<double>(hp, g.n, g.plane, idx, un);
    acc[0] += un * un;
  }
  grid_reduce<1>(acc, rc, scalA + (size_t)blockIdx.y * kScalStride + kOffNrm2);
}

// ------------------------------------------------------------------ K5
// Stoermer-Verlet: u+ = 2u - u_past + tau^2 (L u - m f(u)), evaluated at the clamped site
// (sg_single_sv_solver.hpp:9-19 + boundaries.hpp:40-57); three rotating buffers.
template <int EQ, bool ANISO>
__global__ void __launch_bounds__(kThreads)
k_sv(Geom g, const double *__restrict__ u, const double *__restrict__ upast,
     const double *__restrict__ mf, const double *__restrict__ cf, double *__restrict__ unew,
     double tau2, RedCtx rc, Halo<double> hp, double *slab_scratch) {
  const double *ub = traj(u, g), *up = traj(upast, g);
  double *un = traj(unew, g);
  const double *mb = mf + g.halo;
  const double *cb = ANISO ? cf + g.halo : nullptr;
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    const long long sc = clamp_site(g, idx);
    const double uc = ub[sc];
    const double lu = stencil_site<double, ANISO>(g, ub, cb, sc, 1.0);
    const double buf = lu - mb[sc] * f_real<EQ>(uc);
    const double v = 2 * uc - up[sc] + tau2 * buf;
    un[idx] = v;
    halo_push<double>(hp, g.n, g.plane, idx, v);
  }
  if (rc.slab != nullptr) { // cross-slab barrier: ghost planes complete before the next step reads them
    double acc[1] = {0.0};
    grid_reduce<1>(acc, rc, slab_scratch);
  }
}

// z-slabs: push the first/last owned plane of x into the neighbours' ghost planes, then a
// cross-slab barrier (used after host uploads of stencil inputs: U, C, matfunc/stencil input)
template <class S>
__global__ void __launch_bounds__(kThreads)
k_halo_exchange(Geom g, const S *__restrict__ x, RedCtx rc, Halo<S> hp, double *slab_scratch) {
  const S *xb = x + g.halo;
  for (long long t = (long long)blockIdx.x * kThreads + threadIdx.x; t < 2 * g.plane;
       t += (long long)gridDim.x * kThreads) {
    const long long idx = t < g.plane ? t : g.n - 2 * g.plane + t;
    halo_push<S>(hp, g.n, g.plane, idx, xb[idx]);
  }
  double acc[1] = {0.0};
  grid_reduce<1>(acc, rc, slab_scratch);
}

// v = (u - u_past)/dt     sg_single_driver.cpp:113
__global__ void __launch_bounds__(kThreads)
k_velocity(Geom g, const double *__restrict__ u, const double *__restrict__ upast,
           double *__restrict__ v, double dt) {
  const double *ub = traj(u, g), *up = traj(upast, g);
  double *vb = traj(v, g);
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads)
    vb[idx] = (ub[idx] - up[idx]) / dt;
}

// u_past = u - dt*v       sg_single_driver.cpp:70
__global__ void __launch_bounds__(kThreads)
k_upast_from_v(Geom g, const double *__restrict__ u, const double *__restrict__ v,
               double *__restrict__ upast, double dt) {
  const double *ub = traj(u, g), *vb = traj(v, g);
  double *up = traj(upast, g);
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads)
    up[idx] = ub[idx] - dt * vb[idx];
}

// mass (sum |u|^2) and non-finite count, one launch each
template <class S, int WHAT>
__global__ void __launch_bounds__(kThreads)
k_reduce_state(Geom g, const S *__restrict__ u, double *dst, RedCtx rc) {
  const S *ub = traj(u, g);
  double acc[1] = {0.0};
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    const double a = Sc<S>::abs2(ub[idx]);
    if (WHAT == NLS_RED_MASS) acc[0] += a;
    else acc[0] += isfinite(a) ? 0.0 : 1.0;
  }
  grid_reduce<1>(acc, rc, dst + blockIdx.y);
}

} // namespace nls
