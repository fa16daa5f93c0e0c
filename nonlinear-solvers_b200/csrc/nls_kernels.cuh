// nls_kernels.cuh -- the hot-path kernels (all fp64 SIMT, HBM/L2-bandwidth bound).
//
//   K1  k_stencil            y = L x                         (replaces device/spmv.hpp:65-72)
//   K2  k_lanczos_a/b       one Lanczos iteration in two fused passes over an un-normalised
//                            basis, scalars stay on the device (replaces
//                            device/lanczos_complex.hpp:318-551 and the host algorithm
//                            eigen_krylov_complex.hpp:13-52 / _real.hpp:8-55)
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
//   [193 .. 193+128)    s[i]     : column_i^H w of the current iteration
//   [328 .. 328+128)    g[i]     : column_i^H column_j (Gram entries of the newest column)
//   [456 .. 456+128)    g2[i]    : NLS_REORTH_GRAM: the Gram entries of the newest two columns alternate between g and g2
#pragma once
#include "nls_device.cuh"
#include "../../include/nls.h"

namespace nls {

constexpr int kOffNrm2 = 0;
constexpr int kOffAlpha = 65;
constexpr int kOffH = 193;
constexpr int kOffG = 328;  // g[i] = column_i^H column_j of the current iteration (W doubles each)
constexpr int kOffG2 = 456; // NLS_REORTH_GRAM: the Gram entries of the two newest columns alternate between g and g2
constexpr int kScalStride = 584;
__device__ __host__ __forceinline__ int gram_off(int j) { return (j & 1) ? kOffG2 : kOffG; } // Gram entries of column j
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

// ------------------------------------------------------------------ K2
// Basis layout: column 0 is the start-vector buffer, columns 1..m-1 live in V; columns are kept
// UN-normalised (column i = w_{i-1} as it left the orthogonalisation, v_i = column_i / beta_i,
// beta_i^2 = nrm2[i]), so every column is written exactly once and never rescaled: the 1/beta
// factors are folded into the scalar coefficients.  One Lanczos iteration is two passes:
//   pass A (k_lanczos_a): w = L v_j - beta_j v_{j-1}  (eigen_krylov_complex.hpp:25-27), stored
//           in column j+1, and in the same sweep the sums s_i = column_i^H w and the Gram
//           entries g_i = column_i^H column_j for all i <= j.  T(j,j) = Re s_j / beta_j (:28).
//           The reference subtracts alpha v_j FIRST and takes its re-orthogonalisation dots
//           from w' = w - alpha v_j (:29,:32-33); column_i^H w' = s_i - (alpha/beta_j) g_i gives
//           the same dots without a second sweep (without the g term the alpha component leaks
//           through the existing loss of orthogonality and is amplified by |alpha|/beta per
//           iteration);
//   pass B (k_lanczos_b): w -= alpha v_j + sum_i h_i v_i  (:29,:34), ||w||^2 (:38).
// 2j+5 vector passes per iteration instead of the 2j+10 of "alpha step, then block dots, then
// block update, then normalise" (SURVEY 8d counts the latter).
// NLS_REORTH_GRAM (what NLS_REORTH_FULL runs for m <= 13): pass A with Geom::raw writes r = L column_j and measures only
// s_j, s_{j-1}; the closing pass B (lanczos_b_body<GDR>) derives the older projections from Gram entries: j+6 passes.
template <class S>
__device__ __forceinline__ const S *colptr(const S *start, const S *V, size_t colstride, int i) {
  return i == 0 ? start : V + (size_t)(i - 1) * colstride;
}

// columns [i0, i0+cnt) of 0..j.  FIRST: also forms w (stencil of column j + recurrence term).
template <class S, bool ANISO, int NC, bool FIRST>
__global__ void __launch_bounds__(kThreads, 2)
k_lanczos_a(Geom g, const S *__restrict__ start, S *__restrict__ V, size_t colstride, int j, int i0,
            int cnt, const double *__restrict__ cf, double *scal, RedCtx rc) {
  using T = Sc<S>;
  double *sb = scal + (size_t)blockIdx.y * kScalStride;
  const S *cj = traj(colptr<S>(start, V, colstride, j), g);
  const S *cm = j > 0 ? traj(colptr<S>(start, V, colstride, j - 1), g) : nullptr;
  S *wc = traj(V + (size_t)j * colstride, g); // column j+1
  const S *cb = traj(colptr<S>(start, V, colstride, i0), g);
  const size_t cs0 = i0 == 0 ? (size_t)((V + 0) - start) : colstride; // column i0 -> i0+1 distance
  const double *cfb = ANISO ? cf + g.halo : nullptr;
  double inv = 0.0, cmul = 0.0;
  if (FIRST) {
    const double bj = sqrt(sb[kOffNrm2 + j]);
    inv = 1.0 / bj;
    cmul = j > 0 ? -bj / sqrt(sb[kOffNrm2 + j - 1]) : 0.0; // -beta_j * (1/beta_{j-1})
    if (g.raw) { inv = 1.0; cmul = 0.0; }                   // w = L column_j as it is (NLS_REORTH_GRAM)
  }
  double acc[NC * T::W];
#pragma unroll
  for (int k = 0; k < NC * T::W; ++k) acc[k] = 0.0;
  const long long pfd = (long long)(FIRST ? g.pf_first : g.pf) * gridDim.x * kThreads;
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    if (pfd && idx + pfd < g.n) {
#pragma unroll
      for (int i = 0; i < NC; ++i)
        if (i < cnt) pf_l2((i == 0 ? cb : cb + cs0 + (size_t)(i - 1) * colstride) + idx + pfd);
      if (FIRST) {
        if (j >= i0 + cnt) pf_l2(cj + idx + pfd);
        if (j > 0 && j - 1 >= i0 + cnt) pf_l2(cm + idx + pfd);
        if (ANISO) pf_l2(cfb + idx + pfd);
      } else {
        pf_l2(wc + idx + pfd);
      }
    }
    S v[NC];
#pragma unroll
    for (int i = 0; i < NC; ++i)
      if (i < cnt) v[i] = (i == 0 ? cb : cb + cs0 + (size_t)(i - 1) * colstride)[idx];
    S w;
    if (FIRST) {
      w = stencil_site<S, ANISO>(g, cj, cfb, idx, inv);
      if (j > 0) w = T::fma_r(cmul, cm[idx], w);
      wc[idx] = w;
    } else {
      w = wc[idx];
    }
#pragma unroll
    for (int i = 0; i < NC; ++i)
      if (i < cnt) T::dot_acc(acc + i * T::W, v[i], w);
  }
  grid_reduce<NC * T::W>(acc, rc, sb + kOffH + i0 * T::W);
}

// streaming (evict-first) loads the scheduler may move freely: basis columns are read once per pass
__device__ __forceinline__ double ld_stream(const double *p) {
  double v;
  asm("ld.global.cs.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ double2 ld_stream(const double2 *p) {
  double2 v;
  asm("ld.global.cs.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}

// Pass A, FIRST chunk, 3D grids: plane-marching form (see March in nls_device.cuh).  Same
// arithmetic as k_lanczos_a<.., FIRST = true>; only the order in which sites are visited and
// where the stencil neighbours come from differ.
template <class S, bool ANISO, int NC>
__global__ void __launch_bounds__(kThreads, 2)
k_lanczos_a3(Geom g, March mc, const S *__restrict__ start, S *__restrict__ V, size_t colstride, int j,
             int cnt, const double *__restrict__ cf, double *scal, RedCtx rc) {
  using T = Sc<S>;
  double *sb = scal + (size_t)blockIdx.y * kScalStride;
  const S *cj = traj(colptr<S>(start, V, colstride, j), g);
  const S *cm = j > 0 ? traj(colptr<S>(start, V, colstride, j - 1), g) : nullptr;
  S *wc = traj(V + (size_t)j * colstride, g); // column j+1
  const S *cb = traj(start, g);               // column 0; column i >= 1 at traj(V) + (i-1)*colstride
  const S *vb = traj((const S *)V, g);
  const double *cfb = ANISO ? cf + g.halo : nullptr;
  const double bj = sqrt(sb[kOffNrm2 + j]);
  const double inv = 1.0 / bj;
  const double cmul = j > 0 ? -bj / sqrt(sb[kOffNrm2 + j - 1]) : 0.0;
  double acc[NC * T::W];
#pragma unroll
  for (int k = 0; k < NC * T::W; ++k) acc[k] = 0.0;
  const int tx = threadIdx.x % kTileX, ty = threadIdx.x / kTileX;
  const long long tiles = (long long)mc.ntx * mc.nty;
  for (long long unit = blockIdx.x; unit < mc.units; unit += gridDim.x) {
    const int tile = (int)(unit % tiles), zc = (int)(unit / tiles);
    const int i = (tile % mc.ntx) * kTileX + tx, jj = (tile / mc.ntx) * kTileY + ty;
    if (i >= g.nx || jj >= g.ny) continue;
    const int k0 = zc * mc.zc, k1 = min(k0 + mc.zc, g.nz_loc);
    long long idx = ((long long)k0 * g.ny + jj) * g.nx + i;
    unsigned fxy = 0;
    if (i != 0) fxy |= NB_XM;
    if (i != g.nx - 1) fxy |= NB_XP;
    const bool ring_xy = i == 0 || i == g.nx - 1 || jj == 0 || jj == g.ny - 1;
    // z window (ghost planes of a slab sit at idx -+ plane; the flat-index tests gate the global ends)
    S xc = cj[idx];
    S xzm = (g.z0 + k0 > 0) ? cj[idx - g.plane] : T::zero();
    double cc = 0.0, czm = 0.0;
    if (ANISO) {
      cc = cfb[idx];
      czm = (g.z0 + k0 > 0) ? cfb[idx - g.plane] : 0.0;
    }
    for (int k = k0; k < k1; ++k, idx += g.plane) {
      const int kg = g.z0 + k;
      unsigned f = fxy;
      if (!(kg == 0 && jj == 0)) f |= NB_YM;                        // gi >= nx
      if (!(kg == g.nz_glob - 1 && jj == g.ny - 1)) f |= NB_YP;     // gi + nx < nglob (y-wrap quirk kept)
      if (kg > 0) f |= NB_ZM;
      if (kg < g.nz_glob - 1) f |= NB_ZP;
      if (ring_xy || kg == 0 || kg == g.nz_glob - 1) f |= NB_RING;
      S cmv = T::zero();
      if (j > 0) cmv = ld_stream(cm + idx);
      const S xzp = (f & NB_ZP) ? cj[idx + g.plane] : T::zero();
      const S xym = (f & NB_YM) ? cj[idx - g.nx] : T::zero();
      const S xyp = (f & NB_YP) ? cj[idx + g.nx] : T::zero();
      const S xxm = (f & NB_XM) ? cj[idx - 1] : T::zero();
      const S xxp = (f & NB_XP) ? cj[idx + 1] : T::zero();
      double czp = 0.0, cym = 0.0, cyp = 0.0, cxm = 0.0, cxp = 0.0;
      if (ANISO) {
        czp = (f & NB_ZP) ? cfb[idx + g.plane] : 0.0;
        cym = (f & NB_YM) ? cfb[idx - g.nx] : 0.0;
        cyp = (f & NB_YP) ? cfb[idx + g.nx] : 0.0;
        cxm = (f & NB_XM) ? cfb[idx - 1] : 0.0;
        cxp = (f & NB_XP) ? cfb[idx + 1] : 0.0;
      }
      S w = stencil_vals<S, ANISO>(g, f, inv, xc, xzm, xym, xxm, xxp, xyp, xzp, cc, czm, cym, cxm, cxp, cyp, czp);
      if (j > 0) w = T::fma_r(cmul, cmv, w);
      __stcs(wc + idx, w);
#pragma unroll
      for (int q = 0; q < NC; ++q)
        if (q < cnt) {
          S vq;
          if (q == j) vq = xc;            // already in the z window
          else if (q == j - 1) vq = cmv;  // the recurrence term
          else vq = q == 0 ? ld_stream(cb + idx) : ld_stream(vb + (size_t)(q - 1) * colstride + idx);
          T::dot_acc(acc + q * T::W, vq, w);
        }
      xzm = xc; xc = xzp;
      czm = cc; cc = czp;
    }
  }
  grid_reduce<NC * T::W>(acc, rc, sb + kOffH);
}

// columns [i0, i0+cnt).  LAST: ||w||^2 -> nrm2[j+1], T(j,j) -> alpha[j], halo push of column j+1.
// GDR (NLS_REORTH_GRAM, closing chunk holding ALL columns 0..j): the column slot holds the RAW vector r = L column_j and
// only s_j = column_j^H r and s_{j-1} = column_{j-1}^H r were measured (k_lanczos_a* with Geom::raw).  The update is
//   column_{j+1} = r / beta_j - (beta_j / beta_{j-1}) column_{j-1} - sum_i h_i column_i
// with the projections on the older columns derived from the Gram entries g1 of column_j and g0 of column_{j-1} through
// the recurrence L column_i = beta_i (column_{i+1} - cmul_i column_{i-1} + (alpha_i / beta_i) column_i):
//   column_i^H w = (beta_i / beta_j)(g1_{i+1} + (alpha_i / beta_i) g1_i - cmul_i g1_{i-1}) + cmul_j g0_i      (i <= j-2)
// (tools/prototypes/single_sweep_lanczos.py: gram_delayed; tests/test_single_sweep_prototype.py).
template <class S, int NC, bool LAST, bool GDR = false>
__device__ __forceinline__ void lanczos_b_body(const Geom &g, const S *__restrict__ start, S *__restrict__ V, size_t colstride, int j,
                                               int i0, int cnt, double *scal, const RedCtx &rc, const Halo<S> &hp) {
  using T = Sc<S>;
  double *sb = scal + (size_t)blockIdx.y * kScalStride;
  S *wc = traj(V + (size_t)j * colstride, g);
  const S *cb = traj(colptr<S>(start, V, colstride, i0), g);
  const size_t cs0 = i0 == 0 ? (size_t)((V + 0) - start) : colstride;
  __shared__ S h[NC];
  __shared__ double s_gdr[2]; // GDR: 1 / beta_j, Re s_j
  if (GDR && (int)threadIdx.x < NC) {
    constexpr int W = T::W;
    const int ci = threadIdx.x; // i0 == 0
    const double *g1 = sb + gram_off(j), *g0 = sb + gram_off(j - 1);
    const double nj = sb[kOffNrm2 + j], bj = sqrt(nj), inv = 1.0 / bj;
    const double nm = j > 0 ? sb[kOffNrm2 + j - 1] : 1.0;
    const double cmul = j > 0 ? -bj / sqrt(nm) : 0.0;
    // column_k^H column_j for k <= j
    auto gj = [&](int k) -> S { return k < j ? T::from(g1 + k * W) : T::real(nj); };
    S sj = T::mulr(T::from(sb + kOffH + j * W), inv);
    if (j > 0) sj = T::add(sj, T::mulr(T::conj(T::from(g1 + (j - 1) * W)), cmul));
    S hv = T::zero();
    if (ci < cnt) {
      S si;
      if (ci == j) si = sj;
      else if (ci == j - 1) si = T::add(T::mulr(T::from(sb + kOffH + ci * W), inv), T::real(cmul * nm));
      else {
        const double ni = sb[kOffNrm2 + ci], bi = sqrt(ni);
        S est = T::add(gj(ci + 1), T::mulr(gj(ci), sb[kOffAlpha + ci * W] / bi));
        if (ci > 0) est = T::sub(est, T::mulr(gj(ci - 1), -bi / sqrt(sb[kOffNrm2 + ci - 1])));
        si = T::add(T::mulr(est, bi * inv), T::mulr(T::from(g0 + ci * W), cmul));
      }
      if (ci < j) si = T::sub(si, T::mulr(T::from(g1 + ci * W), T::re(sj) / nj));
      hv = T::mulr(si, 1.0 / sb[kOffNrm2 + ci]);
      if (ci == j - 1) hv = T::sub(hv, T::real(cmul)); // the recurrence term rides on column j-1's coefficient
    }
    h[ci] = hv;
    if (ci == 0) { s_gdr[0] = inv; s_gdr[1] = T::re(sj); }
  }
  if (!GDR && (int)threadIdx.x < NC) { // coefficient of column i: (alpha v_j + h_i v_i) in un-normalised form
    const int i = threadIdx.x;
    S hv = T::zero();
    if (i < cnt) {
      const int ci = i0 + i;
      S si = T::from(sb + kOffH + ci * T::W);
      if (ci < j) { // column_i^H (w - alpha v_j) = s_i - (Re s_j / beta_j^2) g_i
        const double q = sb[kOffH + j * T::W] / sb[kOffNrm2 + j];
        si = T::sub(si, T::mulr(T::from(sb + kOffG + ci * T::W), q));
      }
      hv = T::mulr(si, 1.0 / sb[kOffNrm2 + ci]);
    }
    h[i] = hv;
  }
  __syncthreads();
  // LAST: acc[0 .. NC*W) = g_i = column_i^H column_{j+1} (for the next iteration), acc[NC*W] = ||w||^2
  constexpr int KA = LAST ? NC * T::W + 1 : 1;
  double acc[KA];
#pragma unroll
  for (int k = 0; k < KA; ++k) acc[k] = 0.0;
  const long long pfd = (long long)g.pf * gridDim.x * kThreads;
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    if (pfd && idx + pfd < g.n) {
#pragma unroll
      for (int i = 0; i < NC; ++i)
        if (i < cnt) pf_l2((i == 0 ? cb : cb + cs0 + (size_t)(i - 1) * colstride) + idx + pfd);
      pf_l2(wc + idx + pfd);
    }
    {
      S v[NC];
#pragma unroll
      for (int i = 0; i < NC; ++i)
        if (i < cnt) v[i] = (i == 0 ? cb : cb + cs0 + (size_t)(i - 1) * colstride)[idx];
      S wv = wc[idx];
      if (GDR) wv = T::mulr(wv, s_gdr[0]);
#pragma unroll
      for (int i = 0; i < NC; ++i)
        if (i < cnt) wv = T::sub(wv, T::mul(h[i], v[i]));
      wc[idx] = wv;
      if (LAST) {
#pragma unroll
        for (int i = 0; i < NC; ++i)
          if (i < cnt) T::dot_acc(acc + i * T::W, v[i], wv);
        acc[KA - 1] += T::abs2(wv);
        halo_push<S>(hp, g.n, g.plane, idx, wv); // column j+1 is the next iteration's stencil input
      }
    }
  }
  if (LAST) {
    if (blockIdx.x == 0 && threadIdx.x == 0) // T(j,j) = Re(v_j^H w) = Re s_j / beta_j
      sb[kOffAlpha + j * T::W] = (GDR ? s_gdr[1] : sb[kOffH + j * T::W]) / sqrt(sb[kOffNrm2 + j]);
    grid_reduce2s<KA - 1>(acc, rc, sb + (GDR ? gram_off(j + 1) : kOffG) + i0 * T::W, sb + kOffNrm2 + j + 1);
  }
}

template <class S, int NC>
__global__ void __launch_bounds__(kThreads, 2)
k_lanczos_b_gdr(Geom g, const S *__restrict__ start, S *__restrict__ V, size_t colstride, int j, int i0,
                int cnt, double *scal, RedCtx rc, Halo<S> hp) {
  lanczos_b_body<S, NC, true, true>(g, start, V, colstride, j, i0, cnt, scal, rc, hp);
}

template <class S, int NC, bool LAST>
__global__ void __launch_bounds__(kThreads, 2)
k_lanczos_b(Geom g, const S *__restrict__ start, S *__restrict__ V, size_t colstride, int j, int i0,
            int cnt, double *scal, RedCtx rc, Halo<S> hp) {
  lanczos_b_body<S, NC, LAST>(g, start, V, colstride, j, i0, cnt, scal, rc, hp);
}

// NLS_REORTH_MGS: the reference's sequential sweep, operation for operation (eigen_krylov_complex.hpp:28-38):
//   w -= (v_j^H w) v_j;  for i = 0..j: w -= (v_i^H w) v_i;  ||w||.
// One launch per dot: w -= (H[isub] / nrm2[isub]) column_isub, then H[idot] = column_idot^H w  (NORM: ||w||^2 instead).
// j + 3 launches per iteration instead of 2: a bisecting tool that separates re-orthogonalisation rounding from bugs.
template <class S, bool NORM>
__global__ void __launch_bounds__(kThreads)
k_lanczos_mgs(Geom g, const S *__restrict__ start, S *__restrict__ V, size_t colstride, int j, int isub, int idot,
              int write_alpha, double *scal, RedCtx rc, Halo<S> hp) {
  using T = Sc<S>;
  double *sb = scal + (size_t)blockIdx.y * kScalStride;
  S *wc = traj(V + (size_t)j * colstride, g);
  const S *cs = traj(colptr<S>(start, V, colstride, isub), g);
  const S *cd = NORM ? nullptr : traj(colptr<S>(start, V, colstride, idot), g);
  const S coef = T::mulr(T::from(sb + kOffH + isub * T::W), 1.0 / sb[kOffNrm2 + isub]);
  if (write_alpha && blockIdx.x == 0 && threadIdx.x == 0) // T(j,j) = Re(v_j^H w) = Re s_j / beta_j   (:28)
    sb[kOffAlpha + j * T::W] = sb[kOffH + j * T::W] / sqrt(sb[kOffNrm2 + j]);
  double acc[T::W];
#pragma unroll
  for (int k = 0; k < T::W; ++k) acc[k] = 0.0;
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n; idx += (long long)gridDim.x * kThreads) {
    const S wv = T::sub(wc[idx], T::mul(coef, cs[idx]));
    wc[idx] = wv;
    if (NORM) {
      acc[0] += T::abs2(wv);
      halo_push<S>(hp, g.n, g.plane, idx, wv);
    } else {
      T::dot_acc(acc, cd[idx], wv);
    }
  }
  if (NORM) {
    double a1[1] = {acc[0]};
    grid_reduce<1>(a1, rc, sb + kOffNrm2 + j + 1);
  } else {
    grid_reduce<T::W>(acc, rc, sb + kOffH + idot * T::W);
  }
}

// Gram entries g_i = column_i^H column_j for the columns [i0, i0+cnt) that the LAST chunk of the
// previous pass B did not hold (only when more than NC columns take part, i.e. m > NC + 1)
template <class S, int NC>
__global__ void __launch_bounds__(kThreads)
k_lanczos_g(Geom g, const S *__restrict__ start, const S *__restrict__ V, size_t colstride, int j, int i0,
            int cnt, double *scal, RedCtx rc) {
  using T = Sc<S>;
  double *sb = scal + (size_t)blockIdx.y * kScalStride;
  const S *cj = traj(colptr<S>(start, V, colstride, j), g);
  const S *cb = traj(colptr<S>(start, V, colstride, i0), g);
  const size_t cs0 = i0 == 0 ? (size_t)((V + 0) - start) : colstride;
  double acc[NC * T::W];
#pragma unroll
  for (int k = 0; k < NC * T::W; ++k) acc[k] = 0.0;
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    const S cv = cj[idx];
#pragma unroll
    for (int i = 0; i < NC; ++i)
      if (i < cnt) T::dot_acc(acc + i * T::W, (i == 0 ? cb : cb + cs0 + (size_t)(i - 1) * colstride)[idx], cv);
  }
  grid_reduce<NC * T::W>(acc, rc, sb + kOffG + i0 * T::W);
}

// ------------------------------------------------------------------ K3
// One warp per trajectory: implicit-shift QL on the m x m symmetric tridiagonal T
// (diag = Re alpha_j, alpha_{m-1} = 0 as in the reference; off-diag = sqrt(nrm2[j+1])),
// eigenvector rows distributed over the lanes, then
//   coef_i = beta0 * sum_k Z_ik f(t, lambda_k) Z_0k            (beta * V * f(T) * e1, :77-82)
// with coefficient i divided by beta_i because the basis columns are kept un-normalised.  Function tables: eigen_krylov_complex.hpp:70-76,:89-96,
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

// Warp-cooperative implicit-shift QL (tql2) for m <= 32 with the tridiagonal in REGISTERS: lane i owns d_i and
// e_i (e_i couples i and i+1; e_{m-1} = 0), every lane runs the same scalar recurrence on values broadcast by
// shuffles (fetched one rotation ahead, so no memory round trip sits on the serial rotation chain), the
// deflation scan is one ballot, and lane k owns row k of the eigenvector matrix Z (shared memory, leading
// dimension ldz, must hold the identity on entry).  On exit d (lane k) = lambda_k, column k of Z its eigenvector.
// Any backward-stable eigensolver gives parity here: the result only enters through Q f(L) Q^T e1.
__device__ __forceinline__ bool tridiag_warp_reg(double &d, double &e, double *__restrict__ tz, const int ldz, const int m) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const bool row = lane < m;
  double *zr = tz + (row ? lane : 0) * ldz;
  bool ok = true;
  for (int l = 0; l < m && ok; ++l) {
    int iter = 0;
    for (;;) {
      // smallest mm >= l with a negligible e_mm (bit m-1 is always set: e_{m-1} = 0)
      const double dn = __shfl_down_sync(FULL, d, 1);
      const bool negl = lane >= m - 1 || fabs(e) <= 2.220446049250313e-16 * (fabs(d) + fabs(dn));
      const unsigned mask = __ballot_sync(FULL, negl) & (FULL << l);
      const int mm = __ffs(mask) - 1;
      if (mm == l) break;
      if (iter++ == 60) { ok = false; break; }
      const double dl = __shfl_sync(FULL, d, l), dl1 = __shfl_sync(FULL, d, l + 1), el = __shfl_sync(FULL, e, l);
      const double dmm = __shfl_sync(FULL, d, mm);
      double g = (dl1 - dl) / (2.0 * el);
      double r = sqrt(fma(g, g, 1.0));
      g = dmm - dl + el / (g + copysign(r, g));
      double s = 1.0, c = 1.0, p = 0.0;
      int i = mm - 1;
      bool early = false;
      double e_i = __shfl_sync(FULL, e, i), d_i = __shfl_sync(FULL, d, i), d_hi = dmm;
      double zh = row ? zr[mm] : 0.0;
      for (; i >= l; --i) {
        const int ip = i > l ? i - 1 : i;
        const double e_n = __shfl_sync(FULL, e, ip), d_n = __shfl_sync(FULL, d, ip); // next rotation's pair
        const double zl = row ? zr[i] : 0.0;
        const double f = s * e_i;
        const double b = c * e_i;
        const double t2 = fma(f, f, g * g);
        const double ir = rsqrt(t2);
        r = t2 == 0.0 ? 0.0 : t2 * ir;
        if (lane == i + 1) e = r;
        if (r == 0.0) {
          if (lane == i + 1) d = d_hi - p;
          if (lane == mm) e = 0.0;
          if (row) zr[i + 1] = zh;
          early = true;
          break;
        }
        s = f * ir;
        c = g * ir;
        g = d_hi - p;
        r = (d_i - g) * s + 2.0 * c * b;
        p = s * r;
        if (lane == i + 1) d = g + p;
        g = c * r - b;
        if (row) zr[i + 1] = s * zl + c * zh;
        zh = c * zl - s * zh;
        d_hi = d_i;
        d_i = d_n;
        e_i = e_n;
      }
      if (early) continue;
      if (row) zr[l] = zh;
      if (lane == l) { d -= p; e = g; }
      if (lane == mm) e = 0.0;
    }
  }
  return ok;
}

// coefficient vectors from the decomposition held by tridiag_warp_reg's outputs (m <= 32):
//   co_f[i] = scale_i * sum_k Z_ik f(t_f lambda_k) Z_0k,  scale_i = beta_0 / beta_i (columns are un-normalised)
// zf: 2*32 doubles of shared scratch.  co: [nfn][costride] complex pairs.
__device__ __forceinline__ void tridiag_coefs_reg(const double lam, const double *__restrict__ tz, const int ldz, const int m,
                                                  const TridiagArgs &a, const double scale, double *__restrict__ zf,
                                                  double *__restrict__ co, const int costride) {
  const int lane = threadIdx.x & 31;
  for (int f = 0; f < a.nfn; ++f) {
    __syncwarp();
    if (lane < m) {
      const double2 fv = fn_eval(a.fn[f], a.tr[f], a.ti[f], lam);
      const double z0 = tz[lane]; // Z_0k
      zf[2 * lane] = fv.x * z0;
      zf[2 * lane + 1] = fv.y * z0;
    }
    __syncwarp();
    if (lane < m) {
      const double *zr = tz + lane * ldz;
      double sr = 0.0, si = 0.0;
      for (int k = 0; k < m; ++k) {
        const double z = zr[k];
        sr = fma(z, zf[2 * k], sr);
        si = fma(z, zf[2 * k + 1], si);
      }
      co[f * costride + 2 * lane] = sr * scale;
      co[f * costride + 2 * lane + 1] = si * scale;
    }
  }
  __syncwarp();
}

// One warp per trajectory.  m <= 32: register-resident QL above; larger m: the shared-memory form below.
__global__ void __launch_bounds__(32)
k_tridiag(const double *__restrict__ scal, double *__restrict__ coef, TridiagArgs a, int *status) {
  __shared__ double d[kMaxM], e[kMaxM], Z[kMaxM][kMaxM + 1];
  const int lane = threadIdx.x, m = a.m;
  const double *sb = scal + (size_t)blockIdx.x * kScalStride;
  if (m <= 32) {
    double *tz = &Z[0][0];
    const int ldz = m + 1;
    double dr = lane < m - 1 ? sb[kOffAlpha + lane * a.W] : 0.0;   // T(m-1,m-1) = 0
    double er = lane < m - 1 ? sqrt(sb[kOffNrm2 + lane + 1]) : 0.0;
    if (lane < m)
      for (int k = 0; k < m; ++k) tz[lane * ldz + k] = lane == k ? 1.0 : 0.0;
    __syncwarp();
    if (!tridiag_warp_reg(dr, er, tz, ldz, m) && lane == 0) atomicExch(status, 1);
    __syncwarp();
    const double scale = lane < m ? sqrt(sb[kOffNrm2 + 0]) / sqrt(sb[kOffNrm2 + lane]) : 0.0;
    tridiag_coefs_reg(dr, tz, ldz, m, a, scale, e, coef + (size_t)blockIdx.x * 2 * kCoefStride, kCoefStride);
    return;
  }
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
      const double scale = beta0 / sqrt(sb[kOffNrm2 + i]); // columns are un-normalised
      co[2 * i] = sr * scale;
      co[2 * i + 1] = si * scale;
    }
  }
}

// ------------------------------------------------------------------ K4 helpers
// y(idx) = coef_0 start[idx] + sum_{1<=i<m} coef_i V_{i-1}[idx]   (column 0 is the start buffer)
template <class S>
__device__ __forceinline__ S combine_site(const S *__restrict__ st, const S *__restrict__ vb,
                                          size_t colstride, const double *__restrict__ co, int m,
                                          long long idx);
template <>
__device__ __forceinline__ double combine_site<double>(const double *__restrict__ st,
                                                       const double *__restrict__ vb, size_t colstride,
                                                       const double *__restrict__ co, int m,
                                                       long long idx) {
  double s = st[idx] * co[0];
#pragma unroll 4
  for (int i = 1; i < m; ++i) s = fma(vb[(size_t)(i - 1) * colstride + idx], co[2 * i], s);
  return s;
}
template <>
__device__ __forceinline__ double2 combine_site<double2>(const double2 *__restrict__ st,
                                                         const double2 *__restrict__ vb,
                                                         size_t colstride,
                                                         const double *__restrict__ co, int m,
                                                         long long idx) {
  const double2 v0 = st[idx];
  double2 s = make_double2(v0.x * co[0] - v0.y * co[1], v0.x * co[1] + v0.y * co[0]);
#pragma unroll 4
  for (int i = 1; i < m; ++i) {
    const double2 v = vb[(size_t)(i - 1) * colstride + idx];
    const double cr = co[2 * i], ci = co[2 * i + 1];
    s.x = fma(v.x, cr, fma(-v.y, ci, s.x));
    s.y = fma(v.x, ci, fma(v.y, cr, s.y));
  }
  return s;
}

// out = sum coef_i v_i   (nls_matfunc: beta * V * f(T) * e1)
template <class S>
__global__ void __launch_bounds__(kThreads)
k_combine(Geom g, const S *__restrict__ start, const S *__restrict__ V, size_t colstride,
          const double *__restrict__ coef, int fsel, int m, S *__restrict__ out) {
  __shared__ double co[kCoefStride];
  const double *cg = coef + ((size_t)blockIdx.y * 2 + fsel) * kCoefStride;
  for (int i = threadIdx.x; i < 2 * m; i += kThreads) co[i] = cg[i];
  __syncthreads();
  const S *vb = traj(V, g), *st = traj(start, g);
  S *ob = traj(out, g);
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads)
    ob[idx] = combine_site<S>(st, vb, colstride, co, m, idx);
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
k_nlse_finish(Geom g, const double2 *__restrict__ start, const double2 *__restrict__ V, size_t colstride,
              const double *__restrict__ coef, int m,
              const double2 *__restrict__ uold, const double *__restrict__ mf,
              double2 *__restrict__ uout, double2 *__restrict__ rho, NlseParams p, double *scal,
              RedCtx rc, Halo<double2> hp) {
  __shared__ double co[kCoefStride];
  const double *cg = coef + ((size_t)blockIdx.y * 2) * kCoefStride;
  for (int i = threadIdx.x; i < 2 * m; i += kThreads) co[i] = cg[i];
  __syncthreads();
  const double2 *vb = traj(V, g), *st = traj(start, g), *uo = traj(uold, g);
  double2 *un = traj(uout, g), *rb = traj(rho, g);
  const double *mb = mf + g.halo;
  double acc[1] = {0.0};
  const long long pfd = (long long)g.pf_fin * gridDim.x * kThreads;
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    if (pfd && idx + pfd < g.n) { // the m basis columns of this thread's next sweep -> L2 (pf_l2, nls_device.cuh)
      pf_l2(st + idx + pfd);
      for (int i = 1; i < m; ++i) pf_l2(vb + (size_t)(i - 1) * colstride + idx + pfd);
      if (EQ != NLS_EQ_NLSE_CUBIC) pf_l2(uo + idx + pfd);
    }
    const long long sc = clamp_site(g, idx);
    const double2 buf = combine_site<double2>(st, vb, colstride, co, m, sc);
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

// ------------------------------------------------------------------ sEWI (NLS_INT_SEWI)
// B(u) = -m g(|u|^2) u   (nlse_cubic_gautschi_solver.hpp:27-32, nlse_cubic_quintic_gautschi_solver.hpp:19-26,
// nlse_saturating_gautschi_solver.hpp:21-30: real denominator 1 + kappa |u|^2 here), ||B||^2 -> nrm2[0]
template <int EQ>
__global__ void __launch_bounds__(kThreads)
k_sewi_B(Geom g, const double2 *__restrict__ u, const double *__restrict__ mf, double2 *__restrict__ out,
         NlseParams p, double *scal, RedCtx rc, Halo<double2> hp) {
  const double2 *ub = traj(u, g);
  double2 *ob = traj(out, g);
  const double *mb = mf + g.halo;
  double acc[1] = {0.0};
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    const double2 uv = ub[idx];
    const double r = uv.x * uv.x + uv.y * uv.y;
    double gg;
    if (EQ == NLS_EQ_NLSE_CUBIC) gg = mb[idx] * r;
    else if (EQ == NLS_EQ_NLSE_CQ) gg = mb[idx] * (p.s1 * r + p.s2 * (r * r));
    else gg = mb[idx] * (r * (1.0 / (1.0 + p.kappa * r)));
    const double2 b = make_double2(-gg * uv.x, -gg * uv.y);
    ob[idx] = b;
    halo_push<double2>(hp, g.n, g.plane, idx, b);
    acc[0] += b.x * b.x + b.y * b.y;
  }
  grid_reduce<1>(acc, rc, scal + (size_t)blockIdx.y * kScalStride + kOffNrm2);
}

// out = sum coef_i column_i with ||out||^2 -> nrm2[0]: a matrix-function result that is the start
// vector of the next Lanczos run (phi = sincm_multiply(L, B, dt))
template <class S>
__global__ void __launch_bounds__(kThreads)
k_combine_norm(Geom g, const S *__restrict__ start, const S *__restrict__ V, size_t colstride,
               const double *__restrict__ coef, int m, S *__restrict__ out, double *scal, RedCtx rc,
               Halo<S> hp) {
  __shared__ double co[kCoefStride];
  const double *cg = coef + ((size_t)blockIdx.y * 2) * kCoefStride;
  for (int i = threadIdx.x; i < 2 * m; i += kThreads) co[i] = cg[i];
  __syncthreads();
  const S *vb = traj(V, g), *st = traj(start, g);
  S *ob = traj(out, g);
  double acc[1] = {0.0};
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    const S y = combine_site<S>(st, vb, colstride, co, m, idx);
    ob[idx] = y;
    halo_push<S>(hp, g.n, g.plane, idx, y);
    acc[0] += Sc<S>::abs2(y);
  }
  grid_reduce<1>(acc, rc, scal + (size_t)blockIdx.y * kScalStride + kOffNrm2);
}

// u+ = b - sg * 2 tau * a, b = sum coef_i column_i (expm of u-), tau = i dt, followed by the no-flux
// copy (clamped-site read).  fprime = -sg * 2 * dt.
__global__ void __launch_bounds__(kThreads)
k_sewi_final(Geom g, const double2 *__restrict__ start, const double2 *__restrict__ V, size_t colstride,
             const double *__restrict__ coef, int m, const double2 *__restrict__ a, double2 *__restrict__ unew,
             double fprime) {
  __shared__ double co[kCoefStride];
  const double *cg = coef + ((size_t)blockIdx.y * 2) * kCoefStride;
  for (int i = threadIdx.x; i < 2 * m; i += kThreads) co[i] = cg[i];
  __syncthreads();
  const double2 *vb = traj(V, g), *st = traj(start, g), *ab = traj(a, g);
  double2 *un = traj(unew, g);
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    const long long sc = clamp_site(g, idx);
    const double2 b = combine_site<double2>(st, vb, colstride, co, m, sc);
    const double2 av = ab[sc];
    un[idx] = make_double2(b.x - fprime * av.y, b.y + fprime * av.x); // b + fprime * i * a
  }
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
             const double *__restrict__ coef, int m,
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
  const double *vb = traj(V, g), *ub = traj(u, g);
  double *up = traj(upast, g), *go = traj(gout, g), *ao = traj(accout, g);
  const double *mb = mf + g.halo;
  double acc[1] = {0.0};
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    const double uv = ub[idx]; // column 0 of this Lanczos run is u itself
    double sf = uv * cof[0], sc = uv * coc[0];
#pragma unroll 4
    for (int i = 1; i < m; ++i) {
      const double v = vb[(size_t)(i - 1) * colstride + idx];
      sf = fma(v, cof[2 * i], sf);
      sc = fma(v, coc[2 * i], sc);
    }
    const double gv = -(mb[idx] * f_real<EQ>(sf));
    go[idx] = gv;
    halo_push<double>(hp, g.n, g.plane, idx, gv);
    acc[0] += gv * gv;
    ao[idx] = 2 * sc - up[idx];
    up[idx] = uv;
  }
  grid_reduce<1>(acc, rc, scalB + (size_t)blockIdx.y * kScalStride + kOffNrm2);
}

// Gautschi stage B: u = acc + tau^2 * beta V sinc2(T) e1, no-flux copy, ||u||^2 for the next step
template <int DUMMY>
__global__ void __launch_bounds__(kThreads)
k_gautschi_b(Geom g, const double *__restrict__ start, const double *__restrict__ V, size_t colstride,
             const double *__restrict__ coef, int m,
             const double *__restrict__ accin, double *__restrict__ u, double tau2, double *scalA,
             RedCtx rc, Halo<double> hp) {
  __shared__ double co[kCoefStride];
  const double *cg = coef + ((size_t)blockIdx.y * 2) * kCoefStride;
  for (int i = threadIdx.x; i < 2 * m; i += kThreads) co[i] = cg[i];
  __syncthreads();
  const double *vb = traj(V, g), *st = traj(start, g), *ai = traj(accin, g);
  double *ub = traj(u, g);
  double acc[1] = {0.0};
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    const long long sc = clamp_site(g, idx);
    const double s = combine_site<double>(st, vb, colstride, co, m, sc);
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

// NLS_RED_ENERGY (nls.h): the drift scripts' energies in one pass over u (neighbours through L1/L2).
//   complex: compare_utils_complex_2d.py:134-151, real: energy_sge.py:6-14 (numpy.gradient edges).
// h0/h1/h2: spacing used along x / y / z.  upast, inv_dt: real states only.
template <class S>
__global__ void __launch_bounds__(kThreads)
k_reduce_energy(Geom g, const S *__restrict__ u, const double *__restrict__ upast, double inv_dt, double h0,
                double h1, double h2, double *dst, RedCtx rc) {
  using T = Sc<S>;
  constexpr bool CPLX = T::W == 2;
  const S *ub = traj(u, g);
  const double *up = CPLX ? nullptr : traj(upast, g);
  const long long nx = g.nx, pl = (long long)g.nx * g.ny;
  const int nz = g.dim3 ? g.nz_loc : 1;
  double acc[1] = {0.0};
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n;
       idx += (long long)gridDim.x * kThreads) {
    int i, j, k;
    if (g.n <= 0xffffffffLL) { // 32-bit divisions (the 64-bit ones cost more than the loads)
      const unsigned ui = (unsigned)idx, r = ui / (unsigned)g.nx;
      i = (int)(ui - r * (unsigned)g.nx);
      k = (int)(r / (unsigned)g.ny);
      j = (int)(r - (unsigned)k * (unsigned)g.ny);
    } else {
      const long long r = idx / nx;
      i = (int)(idx % nx); j = (int)(r % g.ny); k = (int)(r / g.ny);
    }
    const S c = ub[idx];
    const double a2 = T::abs2(c);
    if (CPLX) {
      double e = -0.5 * a2 * a2;
      if (i > 0 && i < g.nx - 1 && j > 0 && j < g.ny - 1 && (nz == 1 || (k > 0 && k < nz - 1))) {
        e += T::abs2(T::sub(ub[idx + 1], ub[idx - 1])) / (4.0 * h0 * h0);
        e += T::abs2(T::sub(ub[idx + nx], ub[idx - nx])) / (4.0 * h1 * h1);
        if (nz > 1) e += T::abs2(T::sub(ub[idx + pl], ub[idx - pl])) / (4.0 * h2 * h2);
      }
      acc[0] += e;
    } else {
      // numpy.gradient: centred inside, first-order one-sided on the faces
      const long long xm = i > 0 ? 1 : 0, xp = i < g.nx - 1 ? 1 : 0;
      const long long ym = j > 0 ? nx : 0, yp = j < g.ny - 1 ? nx : 0;
      const double gx = Sc<double>::from((const double *)&ub[idx + xp]) - Sc<double>::from((const double *)&ub[idx - xm]);
      const double gy = Sc<double>::from((const double *)&ub[idx + yp]) - Sc<double>::from((const double *)&ub[idx - ym]);
      const double dgx = gx / (h0 * (double)(xm + xp)), dgy = gy / (h1 * (double)((ym + yp) / nx));
      double e = dgx * dgx + dgy * dgy;
      if (nz > 1) {
        const long long zm = k > 0 ? pl : 0, zp = k < nz - 1 ? pl : 0;
        const double gz = Sc<double>::from((const double *)&ub[idx + zp]) - Sc<double>::from((const double *)&ub[idx - zm]);
        const double dgz = gz / (h2 * (double)((zm + zp) / pl));
        e += dgz * dgz;
      }
      const double uc = Sc<double>::from((const double *)&c);
      const double v = (uc - up[idx]) * inv_dt;
      acc[0] += 0.5 * (e + v * v + 2.0 * (1.0 - cos(uc)));
    }
  }
  grid_reduce<1>(acc, rc, dst + blockIdx.y);
}

// ---- CSR compatibility check of the B2 class layer (include/nls/solvers.hpp): the reference's solver classes are
// handed an assembled CSR (device/spmv.hpp:30-52, sg_single_dev.hpp:28-31).  The engine is matrix-free; a CSR is
// accepted only if it IS the handle's stencil: both operators are applied to a deterministic pseudo-random
// vector and compared entry by entry (any differing matrix entry shows up with probability 1).
__device__ __forceinline__ double hash_unit(unsigned long long i, unsigned long long seed) { // splitmix64 -> (-1, 1)
  unsigned long long z = (i + 1) * 0x9e3779b97f4a7c15ull + seed;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  z ^= z >> 31;
  return (double)(long long)(z >> 11) * (2.0 / 9007199254740992.0) - 1.0;
}
template <class S>
__global__ void __launch_bounds__(kThreads) k_fill_hash(Geom g, S *__restrict__ x, unsigned long long seed) {
  S *xb = traj(x, g);
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n; idx += (long long)gridDim.x * kThreads) {
    const unsigned long long gi = (unsigned long long)(idx + g.goff);
    if (Sc<S>::W == 2) ((double *)(xb + idx))[0] = hash_unit(2 * gi, seed), ((double *)(xb + idx))[1] = hash_unit(2 * gi + 1, seed);
    else ((double *)(xb + idx))[0] = hash_unit(gi, seed);
  }
}
// out[0] = max |(CSR x - y)_i|, out[1] = max |y_i|  (bit patterns of non-negative doubles order like integers)
template <class S, bool VCPLX>
__global__ void __launch_bounds__(kThreads)
k_csr_residual(long long n, const int *__restrict__ rp, const int *__restrict__ ci, const double *__restrict__ val,
               const S *__restrict__ x, const S *__restrict__ y, unsigned long long *out) {
  using T = Sc<S>;
  double dmax = 0.0, ymax = 0.0;
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += (long long)gridDim.x * kThreads) {
    S acc = T::zero();
    for (int q = rp[i]; q < rp[i + 1]; ++q) {
      const int c = ci[q];
      if (c < 0 || c >= n) { dmax = 1e300; continue; }
      const S xv = x[c];
      if (VCPLX) acc = T::add(acc, T::mul(T::from(val + 2 * (size_t)q), xv));
      else acc = T::fma_r(val[q], xv, acc);
    }
    const S d = T::sub(acc, y[i]);
    dmax = fmax(dmax, sqrt(T::abs2(d)));
    ymax = fmax(ymax, sqrt(T::abs2(y[i])));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    dmax = fmax(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
    ymax = fmax(ymax, __shfl_xor_sync(0xffffffffu, ymax, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(out, (unsigned long long)__double_as_longlong(dmax));
    atomicMax(out + 1, (unsigned long long)__double_as_longlong(ymax));
  }
}

// ---- on-device field / initial-condition generators (include/nls.h nls_generate_field) -------------
// Closed-form fields of finalized_scripts/m_fields_2d.py:14-147 and c_fields_2d.py:12-169 (and the bench's analytic
// initial conditions) evaluated at the grid points x_i = linspace(-L, L, n)[i] of the GLOBAL grid, so a z-slab
// handle fills its own planes without any host array.  A field is a sum of terms followed by one post-operation.
__device__ __forceinline__ double gen_coord(int i, int n, double L) { // numpy.linspace(-L, L, n)[i]
  if (n <= 1) return -L;
  const double step = (L - (-L)) / (double)(n - 1);
  return i == n - 1 ? L : (double)i * step + (-L);
}
__device__ __forceinline__ double2 gen_term_eval(const nls_gen_term &t, double x, double y, double z, bool dim3) {
  const double *p = t.p;
  switch (t.kind) {
  case NLS_GEN_CONST: return make_double2(p[0], 0.0);
  case NLS_GEN_GAUSS: { // amp * exp(-r^2 / (2 w^2))      m_fields_2d.py:96-99
    double r2 = (x - p[1]) * (x - p[1]) + (y - p[2]) * (y - p[2]);
    if (dim3) r2 += (z - p[3]) * (z - p[3]);
    return make_double2(p[0] * exp(-r2 / (2 * (p[4] * p[4]))), 0.0);
  }
  case NLS_GEN_COSWAVE: { // amp * cos|sin(f * (k . x) + phase)   m_fields_2d.py:121-125, c_fields_2d.py:60-66
    double kd = p[1] * x + p[2] * y;
    if (dim3) kd += p[3] * z;
    const double a = p[5] * kd + p[4];
    return make_double2(p[0] * (t.iparam ? sin(a) : cos(a)), 0.0);
  }
  case NLS_GEN_TANH_BOUNDARY: { // m1 + (m2 - m1) * 0.5 (1 + tanh(boundary / width))   m_fields_2d.py:17-36
    double b;
    switch (t.iparam) {
    case 0: b = sqrt(x * x + y * y) - p[2]; break;
    case 1: b = fmax(fabs(x), fabs(y)) - p[2]; break;
    case 2: b = x; break; // boundary along the engine's x axis
    case 3: b = y; break;
    default: b = y + x; break;
    }
    return make_double2(p[0] + (p[1] - p[0]) * (0.5 * (1 + tanh(b / p[3]))), 0.0);
  }
  case NLS_GEN_SINSIN: // base * (1 + amp * sin(k x) sin(k y))   c_fields_2d.py:15-19
    return make_double2(p[0] * (1 + p[1] * (sin(p[2] * y) * sin(p[2] * x))), 0.0);
  case NLS_GEN_LAYERS: { // base, or base * contrast inside the odd layers along one axis   c_fields_2d.py:21-33
    const double c = p[3] != 0.0 ? y : x, L = p[2];
    const double lw = 2 * L / (double)t.iparam;
    double v = p[0];
    for (int i = 1; i < t.iparam; i += 2) {
      const double lo = -L + i * lw, hi = -L + (i + 1) * lw;
      if (c >= lo && c < hi) v = p[0] * p[1];
    }
    return make_double2(v, 0.0);
  }
  case NLS_GEN_CHECKER: { // base * tanh(sharp * sin(pi x / cell) sin(pi y / cell))   c_fields_2d.py:36-46
    const double pi = 3.141592653589793;
    const double pat = sin(pi * y / p[1]) * sin(pi * x / p[1]);
    if (p[2] > 0) return make_double2(p[0] * tanh(p[2] * pat), 0.0);
    return make_double2(p[0] * (pat > 0 ? 1.0 : pat < 0 ? -1.0 : 0.0), 0.0);
  }
  case NLS_GEN_HALFSPACE: // base * tanh(sharp * coordinate / L)   c_fields_2d.py:47-48
    return make_double2(p[0] * tanh(p[1] * (p[3] != 0.0 ? y : x) / p[2]), 0.0);
  case NLS_GEN_GAUSS_WAVE: { // amp * exp(-r^2 / (2 s^2)) * exp(i k . x): the Gaussian-bump initial conditions
    double r2 = (x - p[1]) * (x - p[1]) + (y - p[2]) * (y - p[2]);
    double ph = p[5] * x + p[6] * y;
    if (dim3) { r2 += (z - p[3]) * (z - p[3]); ph += p[7] * z; }
    const double a = p[0] * exp(-r2 / (2 * (p[4] * p[4])));
    double sn, cs;
    sincos(ph, &sn, &cs);
    return make_double2(a * cs, a * sn);
  }
  default: { // NLS_GEN_KINK: amp * atan(exp(a - b sqrt((x-y)^2/3 + (x+y)^2/2)))   root sg_driver.cpp:62-64
    const double r = sqrt((x - y) * (x - y) / 3.0 + (x + y) * (x + y) / 2.0);
    return make_double2(p[0] * atan(exp(p[1] - p[2] * r)), 0.0);
  }
  }
}
// pass 1: out = sum of terms (POST_GRADIENT_MASS: from the c field); mm[0..2] = ordered-int min, max, max|.|
template <class S>
__global__ void __launch_bounds__(kThreads)
k_generate(Geom g, const nls_gen_term *__restrict__ terms, int nterms, int post, double p0, double p1, double p2,
           const double *__restrict__ cf, double Lx, double Ly, double Lz, S *__restrict__ out, long long ostride,
           unsigned long long *mm) {
  S *ob = out + (size_t)blockIdx.y * ostride + g.halo;
  double vmin = 1e300, vmax = -1e300, amax = 0.0;
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n; idx += (long long)gridDim.x * kThreads) {
    const unsigned ugi = (unsigned)(idx + g.goff);
    const int i = (int)(ugi % (unsigned)g.nx), jj = (int)(ugi / (unsigned)g.nx);
    const int j = g.dim3 ? jj % g.ny : jj, k = g.dim3 ? jj / g.ny : 0;
    double2 v = make_double2(0.0, 0.0);
    if (post == NLS_GENPOST_GRADIENT_MASS) { // m_fields_2d.py:38-49: central differences / 2 of c, zero on the edges
      const double *c = cf + g.halo;
      const double gy = (j >= 1 && j <= g.ny - 2) ? (c[idx + g.nx] - c[idx - g.nx]) / 2 : 0.0; // script axis 0 = engine y
      const double gx = (i >= 1 && i <= g.nx - 2) ? (c[idx + 1] - c[idx - 1]) / 2 : 0.0;
      const double gs = gy * gy + gx * gx;
      v.x = p0 * (1 + p1 * gs / (gs + p2 * p2));
    } else {
      const double x = gen_coord(i, g.nx, Lx), y = gen_coord(j, g.ny, Ly), z = g.dim3 ? gen_coord(k, g.nz_glob, Lz) : 0.0;
      for (int q = 0; q < nterms; ++q) {
        const double2 t = gen_term_eval(terms[q], x, y, z, g.dim3 != 0);
        v.x += t.x;
        v.y += t.y;
      }
      if (post == NLS_GENPOST_CLAMP_MIN) v.x = fmax(v.x, p0);
    }
    if (Sc<S>::W == 2) ((double *)(ob + idx))[0] = v.x, ((double *)(ob + idx))[1] = v.y;
    else ((double *)(ob + idx))[0] = v.x;
    vmin = fmin(vmin, v.x);
    vmax = fmax(vmax, v.x);
    amax = fmax(amax, Sc<S>::W == 2 ? sqrt(v.x * v.x + v.y * v.y) : fabs(v.x));
  }
  if (post == NLS_GENPOST_MINMAX || post == NLS_GENPOST_MAXABS) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      vmin = fmin(vmin, __shfl_xor_sync(0xffffffffu, vmin, o));
      vmax = fmax(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
      amax = fmax(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    }
    if ((threadIdx.x & 31) == 0) { // order-preserving map of doubles to unsigned integers
      auto key = [](double d) { const long long b = __double_as_longlong(d); return (unsigned long long)(b < 0 ? ~b : (b | (long long)0x8000000000000000ull)); };
      atomicMin(mm + 3 * blockIdx.y, key(vmin));
      atomicMax(mm + 3 * blockIdx.y + 1, key(vmax));
      atomicMax(mm + 3 * blockIdx.y + 2, key(amax));
    }
  }
}
// pass 2 of the normalising post-operations
template <class S>
__global__ void __launch_bounds__(kThreads)
k_generate_post(Geom g, int post, double p0, S *__restrict__ out, long long ostride, const unsigned long long *mm) {
  S *ob = out + (size_t)blockIdx.y * ostride + g.halo;
  auto unkey = [](unsigned long long k) { const long long b = (k & 0x8000000000000000ull) ? (long long)(k & 0x7fffffffffffffffull) : ~(long long)k; return __longlong_as_double(b); };
  const double vmin = unkey(mm[3 * blockIdx.y]), vmax = unkey(mm[3 * blockIdx.y + 1]), amax = unkey(mm[3 * blockIdx.y + 2]);
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < g.n; idx += (long long)gridDim.x * kThreads) {
    double *q = (double *)(ob + idx);
    if (post == NLS_GENPOST_MINMAX) q[0] = p0 * ((q[0] - vmin) / (vmax - vmin)); // c_fields_2d.py:68-72
    else {
      q[0] = q[0] / amax; // u / abs(u).max()
      if (Sc<S>::W == 2) q[1] = q[1] / amax;
    }
  }
}

// Multilinear resampling of a state onto a coarser linspace grid (nls_get_field_resampled):
// downsampling.py:68-104 / :215-258.  One thread per target point; out is dense [batch][tz][ty][tx].
// VEL: sample v = (u - u_past)/dt instead of u (real states).
__device__ __forceinline__ void resample_axis(int a, int t, int n, int &i0, double &f) {
  const double pos = t > 1 ? (double)a * (double)(n - 1) / (double)(t - 1) : 0.0;
  i0 = min((int)pos, n - 2);
  i0 = max(i0, 0);
  f = n > 1 ? pos - (double)i0 : 0.0;
}
template <class S, bool VEL>
__global__ void __launch_bounds__(kThreads)
k_resample(Geom g, const S *__restrict__ u, const double *__restrict__ upast, double inv_dt, int tx, int ty,
           int tz, S *__restrict__ out) {
  using T = Sc<S>;
  const S *ub = traj(u, g);
  const double *up = VEL ? traj(upast, g) : nullptr;
  const int nz = g.dim3 ? g.nz_loc : 1;
  const long long nx = g.nx, pl = (long long)g.nx * g.ny, tot = (long long)tx * ty * tz;
  S *ob = out + (size_t)blockIdx.y * tot;
  auto at = [&](long long p) -> S {
    if (VEL) return T::mulr(T::sub(ub[p], T::from((const double *)&up[p])), inv_dt);
    return ub[p];
  };
  for (long long o = (long long)blockIdx.x * kThreads + threadIdx.x; o < tot; o += (long long)gridDim.x * kThreads) {
    const int a = (int)(o % tx), b = (int)((o / tx) % ty), c = (int)(o / ((long long)tx * ty));
    int i, j, k;
    double fx, fy, fz;
    resample_axis(a, tx, g.nx, i, fx);
    resample_axis(b, ty, g.ny, j, fy);
    resample_axis(c, tz, nz, k, fz);
    const long long dxs = g.nx > 1 ? 1 : 0, dys = g.ny > 1 ? nx : 0, dzs = nz > 1 ? pl : 0;
    const long long p = ((long long)k * g.ny + j) * nx + i;
    auto plane = [&](long long q) -> S {
      const S r0 = T::add(T::mulr(at(q), 1.0 - fx), T::mulr(at(q + dxs), fx));
      const S r1 = T::add(T::mulr(at(q + dys), 1.0 - fx), T::mulr(at(q + dys + dxs), fx));
      return T::add(T::mulr(r0, 1.0 - fy), T::mulr(r1, fy));
    };
    S v = plane(p);
    if (nz > 1) v = T::add(T::mulr(v, 1.0 - fz), T::mulr(plane(p + dzs), fz));
    ob[o] = v;
  }
}

// ---- spectral downsampling (nls_get_field_fft_resampled): the launcher's `downsample_fft`
// (finalized_scripts/downsampling.py:10-34): ortho fft2 -> fftshift -> centred t x t crop -> ifftshift -> ortho ifft2.
// Only t of the N frequencies per axis survive, so the whole pipeline is two small dense products
//   out = Wy (ty x ny) . u (ny x nx) . Wx^T (nx x tx),   W[a][x] = 1/sqrt(N t) sum_kappa e^{2 pi i (kappa a / t - k(kappa) x / N)}
// with k(kappa) the original index of kept frequency kappa: a direct cropped DFT, no FFT library.
__global__ void __launch_bounds__(kThreads) k_dft_weights(double2 *__restrict__ W, int N, int t) {
  const long long tot = (long long)t * N;
  const int start = (N - t) / 2;
  const double scale = 1.0 / sqrt((double)N * (double)t);
  for (long long o = (long long)blockIdx.x * kThreads + threadIdx.x; o < tot; o += (long long)gridDim.x * kThreads) {
    const long long a = o / N, x = o % N;
    double sr = 0.0, si = 0.0;
    for (int kap = 0; kap < t; ++kap) {
      const long long k = (((long long)start + (kap + t / 2) % t - N / 2) % N + N) % N; // fftshift / crop / ifftshift
      // phase in turns, reduced exactly in integers before the division
      const double ph = (double)((kap * a) % t) / (double)t - (double)((k * x) % N) / (double)N;
      double sn, cs;
      sincospi(2.0 * ph, &sn, &cs);
      sr += cs;
      si += sn;
    }
    W[o] = make_double2(sr * scale, si * scale);
  }
}
// T1[b][a][i] = sum_j Wy[a][j] * u[b][j][i]     (S = state scalar; VEL: u -> (u - u_past)/dt)
template <class S, bool VEL>
__global__ void __launch_bounds__(kThreads)
k_dft_cols(Geom g, const S *__restrict__ u, const double *__restrict__ upast, double inv_dt, const double2 *__restrict__ Wy, int ty,
           double2 *__restrict__ T1) {
  constexpr int TA = 4;
  const S *ub = traj(u, g);
  const double *up = VEL ? traj(upast, g) : nullptr;
  const int i = blockIdx.x * kThreads + threadIdx.x;
  const int a0 = blockIdx.z * TA;
  if (i >= g.nx) return;
  double2 acc[TA];
#pragma unroll
  for (int q = 0; q < TA; ++q) acc[q] = make_double2(0.0, 0.0);
  for (int j = 0; j < g.ny; ++j) {
    const long long p = (long long)j * g.nx + i;
    double2 v;
    if (Sc<S>::W == 2) v = make_double2(((const double *)(ub + p))[0], ((const double *)(ub + p))[1]);
    else v = make_double2(VEL ? (((const double *)(ub + p))[0] - up[p]) * inv_dt : ((const double *)(ub + p))[0], 0.0);
#pragma unroll
    for (int q = 0; q < TA; ++q)
      if (a0 + q < ty) {
        const double2 w = Wy[(size_t)(a0 + q) * g.ny + j];
        acc[q].x = fma(w.x, v.x, fma(-w.y, v.y, acc[q].x));
        acc[q].y = fma(w.x, v.y, fma(w.y, v.x, acc[q].y));
      }
  }
#pragma unroll
  for (int q = 0; q < TA; ++q)
    if (a0 + q < ty) T1[((size_t)blockIdx.y * ty + a0 + q) * g.nx + i] = acc[q];
}
// out[b][a][c] = sum_i T1[b][a][i] * Wx[c][i]   (real states keep the real part, like the script)
template <class S>
__global__ void __launch_bounds__(kThreads)
k_dft_rows(const double2 *__restrict__ T1, const double2 *__restrict__ Wx, int nx, int tx, int ty, S *__restrict__ out) {
  const long long tot = (long long)tx * ty;
  for (long long o = (long long)blockIdx.x * kThreads + threadIdx.x; o < tot; o += (long long)gridDim.x * kThreads) {
    const int a = (int)(o / tx), c = (int)(o % tx);
    const double2 *t1 = T1 + ((size_t)blockIdx.y * ty + a) * nx, *w = Wx + (size_t)c * nx;
    double sr = 0.0, si = 0.0;
    for (int i = 0; i < nx; ++i) {
      const double2 tv = t1[i], wv = w[i];
      sr = fma(tv.x, wv.x, fma(-tv.y, wv.y, sr));
      si = fma(tv.x, wv.y, fma(tv.y, wv.x, si));
    }
    double *q = (double *)(out + (size_t)blockIdx.y * tot + o);
    q[0] = sr;
    if (Sc<S>::W == 2) q[1] = si;
  }
}

} // namespace nls
