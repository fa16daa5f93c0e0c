// nls_persist.cuh -- the persistent cooperative time-stepping kernel ("pk").
//
// One launch advances a trajectory by `nsteps` whole time steps.  One CTA per SM, each CTA owns
// one contiguous chunk of the flat site index for the entire launch:
//   * the working vector of the Lanczos process (w) and the stencil input tile with its +-nx
//     halo rows live in shared memory across phases (never re-read from HBM),
//   * basis columns are kept UN-normalised in global memory (column i = w_i, v_i = w_i/beta_i;
//     the 1/beta_i factors are folded into the scalar coefficients), so a column is written
//     exactly once and a chunk's columns are only ever re-read by the CTA that wrote them,
//   * the two global reductions of a Lanczos iteration (h = V^H w, ||w||^2) are grid barriers
//     through one L2 counter; every CTA sums the per-CTA partials in the same fixed order, so
//     all CTAs hold bit-identical scalars and T never leaves the chip,
//   * the m x m tridiagonal eigensolve runs redundantly in warp 0 of every CTA (no broadcast),
//   * the basis recombination, the pointwise nonlinear step, the no-flux boundary copy and the
//     next step's start vector + norm are the tail phase of the same kernel.
// Reference semantics: nlsolvers/host/include/eigen_krylov_{complex,real}.hpp (Lanczos and the
// function tables), nlse_*_solver.hpp / sg_single_solver.hpp / kg_solver.hpp (steps),
// boundaries.hpp:40-57 (boundary copy).  Same results as the multi-kernel path of
// nls_kernels.cuh up to re-orthogonalisation rounding (one block Gram-Schmidt pass including
// the alpha term, instead of alpha step + block pass).
#pragma once
#include <climits>
#include "nls_kernels.cuh"

namespace nls {

// Address-space hints (experimental, -DPK_SHARED_HINTS / -DPK_GLOBAL_HINTS): with them every access through
// the context compiles to LDS/STS instead of generic LD/ST (1449 LDS vs 1402 LD in k_pk_gautschi), but on
// B200 that bought nothing measurable (C1 0.1061 vs 0.1057 ms/step, C2 0.571 vs 0.575) and the shared hints
// miscompile k_pk_gautschi (illegal address; k_pk_nlse passes) with nvcc 12.9 -- off by default.
#define PK_HINT_A 1
#define PK_HINT_B 1
#if defined(PK_HINT_A) || defined(PK_HINT_B) || defined(PK_HINT_G)
#define PK_LSH(p) __builtin_assume(__isShared(p))
#define PK_LGL(p) __builtin_assume(__isGlobal(p))
#endif
#ifdef PK_SHARED_HINTS
#define PK_SHARED(p) __builtin_assume(__isShared(p))
#else
#define PK_SHARED(p) do { } while (0)
#endif
#ifdef PK_GLOBAL_HINTS
#define PK_GLOBAL(p) __builtin_assume(__isGlobal(p))
#else
#define PK_GLOBAL(p) do { } while (0)
#endif

#ifndef PK_THREADS
#define PK_THREADS 512
#endif
#ifndef PK_NV
#define PK_NV 8
#endif
constexpr int kPkThreads = PK_THREADS;
constexpr int kPkWarps = kPkThreads / 32;
constexpr int kPkMaxK = 2 * kMaxM + 2;   // doubles per reduction (complex h for up to kMaxM columns)
constexpr int kPkNV = PK_NV;             // columns per unrolled group
constexpr int kPkPartStride = 160;       // >= max CTAs, padded

struct PkArgs {
  Geom g;
  int m, eq, integ, aniso, nsteps, ncta;
  int debug_fault;     // NLS_FLAG_DEBUG_PK_FAULT: the last CTA exits before its first barrier (abort-path test)
  long long csz;       // sites per chunk (chunk c = sites [c*csz, min((c+1)*csz, n)))
  double dt, s1, s2, kappa;
  void *V;             // m columns; column i at V + i*colstride (scalars)
  size_t colstride;
  void *u0, *u1;       // NLSE: u ping-pong (u0 current at entry); real: u0 = u, u1 = u_past
  void *acc;           // Gautschi accumulator 2 cos(.)u - u_past
  const double *mf, *cf;
  double *part;        // [2][kPkMaxK][kPkPartStride] per-CTA partial sums
  unsigned *bar;       // [0] arrival counter, [1] abort flag   (zeroed before every launch)
  int *status;
};

// Phase timing trace (debug builds only, -DPK_TRACE): thread 0 of a CTA stamps clock64() at phase
// boundaries into shared memory; CTAs 0, ncta/2 and ncta-1 print the deltas at kernel exit.
#ifdef PK_TRACE
} // namespace nls
#include <cstdio>
namespace nls {
__shared__ long long pk_tr_t[1024];
__shared__ int pk_tr_tag[1024];
__shared__ int pk_tr_n;
#define PKT(tag)                                                                  \
  do {                                                                            \
    if (threadIdx.x == 0 && pk_tr_n < 1024) {                                     \
      pk_tr_t[pk_tr_n] = clock64();                                               \
      pk_tr_tag[pk_tr_n++] = (tag);                                               \
    }                                                                             \
  } while (0)
#define PKT_INIT()                                                                \
  do {                                                                            \
    if (threadIdx.x == 0) pk_tr_n = 0;                                            \
    __syncthreads();                                                              \
  } while (0)
#define PKT_DUMP(ncta)                                                            \
  do {                                                                            \
    if (threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == (ncta) / 2 || blockIdx.x == (ncta)-1))  \
      for (int q = 1; q < pk_tr_n; ++q)                                           \
        printf("PKT %d %d %d %lld\n", blockIdx.x, q, pk_tr_tag[q], pk_tr_t[q] - pk_tr_t[q - 1]);     \
  } while (0)
#else
#define PKT(tag) do { } while (0)
#define PKT_INIT() do { } while (0)
#define PKT_DUMP(ncta) do { } while (0)
#endif

// ---------------------------------------------------------------- grid barrier
// Grid-wide reduction = per-CTA partial sums in global memory, ONE counter barrier (red.release / ld.acquire spin by
// thread 0), then every CTA sums all partials in the same fixed order.  Measured against the alternatives on B200
// (tools/microbench/allreduce_bench.cu, profiles/r02d_allreduce_microbench.log; 148 CTAs x 512 threads, K = 1 / 10
// doubles): this scheme 2.2 / 4.1 us per round, cooperative-groups grid.sync + gather 2.4 / 4.1, arrival counter
// spread over 8 words 2.3 / 4.2, sentinel-polled partials (barrier folded into the gather: no counter, no second
// round trip) 3.6 / 7.6 (3.0 / 5.8 with one 32-byte sector per entry; 9.6 / 25.8 when every warp polls), two-level
// reduction through one CTA 4.2 / 6.4: polling the partial array itself hot-spots a few L2 sectors with ~150 x
// more requests than one counter word, so the counter stays.
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u64(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

struct GridSync {
  unsigned *ctr;
  unsigned target; // arrivals expected at the next barrier
  unsigned ncta;
  int slot;        // parity of the barrier count: which partial-sum slot is live
  int *status;     // device status word of the handle: 4 = a grid barrier timed out / was aborted
};

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_add_u32(unsigned *p, unsigned v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// All CTAs are co-resident (cooperative launch).  Returns false if the barrier timed out or
// another CTA raised the abort flag (the kernel then unwinds instead of hanging the GPU).
__device__ __forceinline__ bool grid_barrier(GridSync &gs) {
  __shared__ int s_ok;
  __syncthreads();
  if (threadIdx.x == 0) {
    gs.target += gs.ncta;
    red_release_add_u32(gs.ctr, 1u);
    int ok = 1;
    const long long t0 = clock64();
    while (ld_acquire_u32(gs.ctr) < gs.target) {
      if (clock64() - t0 > 4000000000LL || ld_acquire_u32(gs.ctr + 1) != 0u) { // ~2 s
        atomicExch(gs.ctr + 1, 1u);
        atomicExch(gs.status, 4); // reported by nls_sync / nls_get_field: the state is NOT advanced consistently
        ok = 0;
        break;
      }
    }
    s_ok = ok;
  } else {
    gs.target += gs.ncta;
  }
  gs.slot ^= 1;
  __syncthreads();
  return s_ok != 0;
}

// CTA-level partial sums: acc[K] per thread -> one value per CTA in part[slot][k][cta].
// Must be followed by grid_barrier and then pk_gather.
template <int K>
__device__ __forceinline__ void pk_publish(double (&acc)[K], double *part, int slot, double (*s_red)[kPkWarps], int koff = 0) {
  PK_SHARED(s_red);
  PK_GLOBAL(part);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double v = acc[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) s_red[k][warp] = v;
  }
  __syncthreads();
  if (threadIdx.x < K) {
    double v = 0.0;
#pragma unroll
    for (int w = 0; w < kPkWarps; ++w) v += s_red[threadIdx.x][w];
    part[((size_t)slot * kPkMaxK + koff + threadIdx.x) * kPkPartStride + blockIdx.x] = v;
  }
  __syncthreads(); // s_red may be reused by the next publish
}

// after the barrier: every CTA sums the partials of all CTAs in the same order -> s_out[k].
// The (<= 5) loads of a lane are issued together: one L2 round trip per value, not one per load.
__device__ __forceinline__ void pk_gather_nosync(int K, const double *part, int slot, int ncta, double *s_out) {
  PK_SHARED(s_out);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int kPer = (kPkPartStride + 31) / 32;
  for (int k = warp; k < K; k += kPkWarps) {
    const double *p = part + ((size_t)slot * kPkMaxK + k) * kPkPartStride;
    double t[kPer];
#pragma unroll
    for (int q = 0; q < kPer; ++q) {
      const int c = lane + 32 * q;
      t[q] = c < ncta ? __ldcg(p + c) : 0.0;
    }
    double v = 0.0;
#pragma unroll
    for (int q = 0; q < kPer; ++q)
      if (lane + 32 * q < ncta) v += t[q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_out[k] = v;
  }
}
__device__ __forceinline__ void pk_gather(int K, const double *part, int slot, int ncta, double *s_out) {
  pk_gather_nosync(K, part, slot, ncta, s_out);
  __syncthreads();
}

// ---------------------------------------------------------------- in-kernel tridiagonal solve
// One warp.  Implicit-shift QL on (d, e); the rotation scalars are computed redundantly by all
// lanes from a lane-private copy of d/e (shared memory, [index][lane] layout -> conflict-free,
// no warp synchronisation), eigenvector rows are distributed over the lanes.
// Output: co[2*i], co[2*i+1] = (beta0 * Q f(t Lambda) Q^T e1)_i * colscale_i for nfn functions.
struct PkTri {
  double *d;  // [m][32]   lane-private copies
  double *e;  // [m][32]
  double *Z;  // [m][ldz]  eigenvector rows
  int ldz;
};

__device__ __forceinline__ bool pk_tridiag_warp(PkTri &T, int m) {
  const int lane = threadIdx.x & 31;
  double *const td = T.d, *const te = T.e, *const tz = T.Z;
  PK_SHARED(td);
  PK_SHARED(te);
  PK_SHARED(tz);
  const int ldz = T.ldz;
#define D(i) td[(i) * 32 + lane]
#define E(i) te[(i) * 32 + lane]
  bool ok = true;
  for (int l = 0; l < m && ok; ++l) {
    int iter = 0, mm;
    do {
      for (mm = l; mm < m - 1; ++mm) {
        const double dd = fabs(D(mm)) + fabs(D(mm + 1));
        if (fabs(E(mm)) <= 2.220446049250313e-16 * dd) break;
      }
      if (mm != l) {
        if (iter++ == 60) { ok = false; break; }
        double g = (D(l + 1) - D(l)) / (2.0 * E(l));
        double r = sqrt(fma(g, g, 1.0));
        g = D(mm) - D(l) + E(l) / (g + copysign(r, g));
        double s = 1.0, c = 1.0, p = 0.0;
        int i = mm - 1;
        bool early = false;
        // Serial rotation chain with everything it needs next already in registers: D(i+1) of the next
        // rotation is this one's D(i), the next E/D pair is fetched a rotation ahead, and column i+1 of a
        // lane's eigenvector rows is carried in a register (each rotation loads one new column, stores one
        // finished column) -- no shared-memory round trip sits on the dependency chain.
        const int k0 = lane, k1 = lane + 32;
        double e_i = E(i), d_i = D(i), d_hi = D(mm);
        double zh0 = k0 < m ? tz[k0 * ldz + mm] : 0.0, zh1 = k1 < m ? tz[k1 * ldz + mm] : 0.0;
        for (; i >= l; --i) {
          double e_n = 0.0, d_n = 0.0;
          if (i > l) {
            e_n = E(i - 1);
            d_n = D(i - 1);
          }
          const double zl0 = k0 < m ? tz[k0 * ldz + i] : 0.0, zl1 = k1 < m ? tz[k1 * ldz + i] : 0.0;
          const double f = s * e_i;
          const double b = c * e_i;
          const double t2 = fma(f, f, g * g);
          const double ir = rsqrt(t2); // one reciprocal square root instead of sqrt + divide
          r = t2 == 0.0 ? 0.0 : t2 * ir;
          E(i + 1) = r;
          if (r == 0.0) {
            D(i + 1) = d_hi - p;
            E(mm) = 0.0;
            if (k0 < m) tz[k0 * ldz + i + 1] = zh0;
            if (k1 < m) tz[k1 * ldz + i + 1] = zh1;
            early = true;
            break;
          }
          s = f * ir;
          c = g * ir;
          g = d_hi - p;
          r = (d_i - g) * s + 2.0 * c * b;
          p = s * r;
          D(i + 1) = g + p;
          g = c * r - b;
          if (k0 < m) tz[k0 * ldz + i + 1] = s * zl0 + c * zh0;
          if (k1 < m) tz[k1 * ldz + i + 1] = s * zl1 + c * zh1;
          zh0 = c * zl0 - s * zh0;
          zh1 = c * zl1 - s * zh1;
          d_hi = d_i;
          d_i = d_n;
          e_i = e_n;
        }
        if (!early) {
          if (k0 < m) tz[k0 * ldz + l] = zh0;
          if (k1 < m) tz[k1 * ldz + l] = zh1;
        }
        if (early) continue;
        D(l) -= p;
        E(l) = g;
        E(mm) = 0.0;
      }
    } while (mm != l);
  }
#undef D
#undef E
  return ok;
}


// ---------------------------------------------------------------- shared-memory context
// per-site neighbour flags, computed once per launch (the index divisions leave the hot loops)
enum { PF_XM = 1, PF_XP = 2, PF_YM = 4, PF_YP = 8, PF_ZM = 16, PF_ZP = 32, PF_RING = 64 };

// Shared-memory pointers are handed out through accessors (hook for the address-space hints above).
#define PK_ACC(type, name)                                   \
  type name##_;                                              \
  __device__ __forceinline__ type name() const {             \
    type p = name##_;                                        \
    PK_SHARED(p);                                            \
    return p;                                                \
  }
template <class S> struct PkCtx {
  PK_ACC(S *, X)                // stencil tile: sites [s0 - nx, s0 + csz + nx)
  PK_ACC(S *, Wt)               // w tile: sites [s0, s0 + csz)
  PK_ACC(unsigned char *, flag) // [csz] PF_* bits of the owned sites
  PkTri tri;
  typedef double (*RedPtr)[kPkWarps];
  PK_ACC(RedPtr, red)           // [2*kPkNV + 4][kPkWarps]
  PK_ACC(double *, vals)        // [kPkMaxK] gathered sums
  PK_ACC(double *, alpha)       // [kMaxM]   T(j,j)
  PK_ACC(double *, nrm2)        // [kMaxM+1] nrm2[0] = ||start||^2, nrm2[j+1] = ||w_j||^2
  PK_ACC(double *, coef)        // [2][2*kMaxM] complex coefficients (two function sets)
  PK_ACC(S *, gco)              // [kMaxM] re-orthogonalisation coefficients of the current iteration
  long long s0;                 // first owned site
  int cnt;                      // number of owned sites (may be 0)
};

// column i of the Krylov basis: i == 0 is the start-vector buffer, i >= 1 lives in V
template <class S>
__device__ __forceinline__ S *pk_col(const PkArgs &a, S *start, int i) {
  return i == 0 ? start : (S *)a.V + (size_t)(i - 1) * a.colstride;
}

__device__ __forceinline__ unsigned char pk_site_flags(const Geom &g, long long idx) {
  const long long gi = idx + g.goff;
  const unsigned ugi = (unsigned)gi;
  const unsigned i = ugi % (unsigned)g.nx;
  const unsigned jj = ugi / (unsigned)g.nx;
  unsigned f = 0;
  if (i != 0) f |= PF_XM;                      // laplacians.hpp:33
  if (i != (unsigned)(g.nx - 1)) f |= PF_XP;
  if (gi >= g.nx) f |= PF_YM;                  // flat-index test, no plane check (:39, :140)
  if (gi + g.nx < g.nglob) f |= PF_YP;
  bool ring;
  if (g.dim3) {
    if (gi >= g.plane) f |= PF_ZM;             // :145
    if (gi + g.plane < g.nglob) f |= PF_ZP;
    const unsigned j = jj % (unsigned)g.ny, k = jj / (unsigned)g.ny;
    ring = i == 0 || i == (unsigned)(g.nx - 1) || j == 0 || j == (unsigned)(g.ny - 1) || k == 0 ||
           k == (unsigned)(g.nz_glob - 1);
  } else {
    ring = i == 0 || i == (unsigned)(g.nx - 1) || jj == 0 || jj == (unsigned)(g.ny - 1);
  }
  if (ring) f |= PF_RING;
  return (unsigned char)f;
}

// (L x)(site li) from the X tile (x/y neighbours) and the global column (z neighbours).
// Same rows as stencil_site_f (laplacians.hpp:9-218), neighbours gated by the cached flags.
template <class S, bool ANISO>
__device__ __forceinline__ S pk_stencil(const Geom &g, const S *__restrict__ xc, const S *__restrict__ colg,
                                        const double *__restrict__ cfg, unsigned f, int li) {
  using T = Sc<S>;
  PK_SHARED(xc);
  const int nx = g.nx;
  const S z = T::zero();
  if (!ANISO && !g.dim3 && (f & (PF_XM | PF_XP | PF_YM | PF_YP | PF_RING)) == (PF_XM | PF_XP | PF_YM | PF_YP)) {
    // interior 2D site (94% of the warps are uniformly here): the same fma chain as below without the
    // neighbour selects; the z terms of the general chain add exact zeros, so the bits are identical
    S acc = T::fma_r(g.sc, xc[li - nx], z);
    acc = T::fma_r(g.sc, xc[li - 1], acc);
    acc = T::fma_r(-4.0 * g.sc, xc[li], acc);
    acc = T::fma_r(g.sc, xc[li + 1], acc);
    acc = T::fma_r(g.sc, xc[li + nx], acc);
    return acc;
  }
  const S xm = (f & PF_XM) ? xc[li - 1] : z;
  const S xp = (f & PF_XP) ? xc[li + 1] : z;
  const S ym = (f & PF_YM) ? xc[li - nx] : z;
  const S yp = (f & PF_YP) ? xc[li + nx] : z;
  S zm = z, zp = z;
  if (g.dim3) {
    if (f & PF_ZM) zm = __ldcg(colg + li - g.plane);
    if (f & PF_ZP) zp = __ldcg(colg + li + g.plane);
  }
  S acc = z;
  if (!ANISO) {
    const double dg = g.dim3 ? ((f & PF_RING) ? -5.0 : -6.0) : ((f & PF_RING) ? -3.0 : -4.0);
    acc = T::fma_r(g.sc, zm, acc);
    acc = T::fma_r(g.sc, ym, acc);
    acc = T::fma_r(g.sc, xm, acc);
    acc = T::fma_r(dg * g.sc, xc[li], acc);
    acc = T::fma_r(g.sc, xp, acc);
    acc = T::fma_r(g.sc, yp, acc);
    acc = T::fma_r(g.sc, zp, acc);
  } else {
    const double cc = cfg[li];
    const double wxm = (f & PF_XM) ? (cfg[li - 1] + cc) / 2.0 : 0.0;
    const double wxp = (f & PF_XP) ? (cc + cfg[li + 1]) / 2.0 : 0.0;
    const double wym = (f & PF_YM) ? (cfg[li - nx] + cc) / 2.0 : 0.0;
    const double wyp = (f & PF_YP) ? (cc + cfg[li + nx]) / 2.0 : 0.0;
    const double wzm = (f & PF_ZM) ? (cfg[li - g.plane] + cc) / 2.0 : 0.0;
    const double wzp = (f & PF_ZP) ? (cc + cfg[li + g.plane]) / 2.0 : 0.0;
    double ds = 0.0;
    if (f & PF_XM) ds += wxm;
    if (f & PF_XP) ds += wxp;
    if (f & PF_YM) ds += wym;
    if (f & PF_YP) ds += wyp;
    if (f & PF_ZM) ds += wzm;
    if (f & PF_ZP) ds += wzp;
    acc = T::fma_r(wzm * g.sc, zm, acc);
    acc = T::fma_r(wym * g.sc, ym, acc);
    acc = T::fma_r(wxm * g.sc, xm, acc);
    acc = T::fma_r(-ds * g.sc, xc[li], acc);
    acc = T::fma_r(wxp * g.sc, xp, acc);
    acc = T::fma_r(wyp * g.sc, yp, acc);
    acc = T::fma_r(wzp * g.sc, zp, acc);
  }
  return acc;
}

// load the +-nx halo rows of column `col` into the X tile (after a grid barrier).  Up to 4 elements
// per thread are loaded before the first shared-memory store (one L2 round trip for nx <= 1200).
template <class S>
__device__ __forceinline__ void pk_load_halo_part(const PkArgs &a, PkCtx<S> &c, const S *col, int tid, int nth) {
  const int nx = a.g.nx;
  constexpr int kQ = 5;
  if (c.cnt > 0) {
    for (int t0 = tid; t0 < 2 * nx; t0 += kQ * nth) {
      S v[kQ];
      int xi[kQ];
#pragma unroll
      for (int q = 0; q < kQ; ++q) {
        const int t = t0 + q * nth;
        const bool up = t >= nx;
        const long long idx = up ? c.s0 + c.cnt + (t - nx) : c.s0 - nx + t;
        xi[q] = -1;
        if (t < 2 * nx && idx >= 0 && idx < a.g.n) {
          xi[q] = up ? nx + c.cnt + (t - nx) : t;
          v[q] = __ldcg(col + idx);
        }
      }
#pragma unroll
      for (int q = 0; q < kQ; ++q)
        if (xi[q] >= 0) c.X()[xi[q]] = v[q];
    }
  }
}
template <class S>
__device__ __forceinline__ void pk_load_halo_nosync(const PkArgs &a, PkCtx<S> &c, const S *col) {
  pk_load_halo_part<S>(a, c, col, threadIdx.x, kPkThreads);
}
template <class S>
__device__ __forceinline__ void pk_load_halo(const PkArgs &a, PkCtx<S> &c, const S *col) {
  pk_load_halo_nosync<S>(a, c, col);
  __syncthreads();
}

#ifndef PK_COMB_UNROLL
#define PK_COMB_UNROLL 2
#endif
constexpr int kPkCoefRegs = 12;           // recombination keeps the coefficients in registers for m <= 12
constexpr int kPkCombUnroll = PK_COMB_UNROLL; // sites per unrolled group in the recombination loops of k_pk_gautschi
// unroll factor of the phase loops (the compiler interleaves the loads of the unrolled sites)
__host__ __device__ constexpr int pk_unroll(int nc, int w) {
#ifdef PK_UNROLL_TUNE
  return w == 1 ? (nc <= 1 ? 8 : nc <= 4 ? 4 : 2) : (nc <= 1 ? 4 : nc <= 3 ? 2 : 1);
#else
  return w == 1 ? 2 : 1;
#endif
}

// ---------------------------------------------------------------- Lanczos phases
// Column roles in iteration j: "others" = columns 0..j-2 (groups of <= kPkNV, pointer per
// column), column j-1 (recurrence term + dot, loaded once), column j (the X tile).
// Phase A, group of NC other columns starting at i0.  FIRST: also forms w (stencil + recurrence),
// stores it to the Wt tile and accumulates the j-1 and j dots.
// The site loops live in functions whose pointer parameters are __restrict__ and whose scalars are passed
// by value: inside the kernels the context struct sits in local memory and every store through a generic
// pointer forced the compiler to reload it (loop bound included) and to keep the next site's loads behind the
// current site's stores -- one full memory round trip per site (profiles/r01f_pk_gautschi_phase_trace.txt).
// column k of the basis for a chunk: k == 0 -> startp, k >= 1 -> vbase + (k-1)*cs
template <class S>
__device__ __forceinline__ const S *pk_colp(const S *startp, const S *vbase, size_t cs, int k) {
  return k == 0 ? startp : vbase + (size_t)(k - 1) * cs;
}

template <class S, bool ANISO, int NC, bool FIRST>
__device__ __forceinline__ void pk_loop_a(const Geom &g, const int cnt, const int j, const int i0,
                                          const S *__restrict__ xc, S *__restrict__ Wt,
                                          const unsigned char *__restrict__ flag, const S *__restrict__ startp,
                                          const S *__restrict__ vbase, const size_t cs,
                                          const double *__restrict__ cfg, const double inv_bj, const double cm,
                                          double *__restrict__ acc, double *__restrict__ accM,
                                          double *__restrict__ accA) {
#ifdef PK_HINT_A
  PK_LSH(xc); PK_LSH(Wt); PK_LSH(flag); PK_LGL(startp); PK_LGL(vbase);
#endif
  using T = Sc<S>;
  constexpr int W = T::W;
  constexpr int NA = NC > 0 ? NC : 1;
  const S *p[NA];
#pragma unroll
  for (int i = 0; i < NC; ++i) p[i] = pk_colp<S>(startp, vbase, cs, i0 + i);
  const S *colj = pk_colp<S>(startp, vbase, cs, j);
  const S *colm = j > 0 ? pk_colp<S>(startp, vbase, cs, j - 1) : colj;
  double la[NA * W], lm[W], lj[W];
#pragma unroll
  for (int k = 0; k < NA * W; ++k) la[k] = 0.0;
#pragma unroll
  for (int k = 0; k < W; ++k) lm[k] = lj[k] = 0.0;
#pragma unroll(pk_unroll(NC, W))
  for (int li = threadIdx.x; li < cnt; li += kPkThreads) {
    S v[NA];
#pragma unroll
    for (int i = 0; i < NC; ++i) v[i] = __ldcg(p[i] + li);
    S w;
    if (FIRST) {
      S vm = T::zero();
      if (j > 0) vm = __ldcg(colm + li);
      w = T::mulr(pk_stencil<S, ANISO>(g, xc, colj, cfg, flag[li], li), inv_bj);
      w = T::fma_r(cm, vm, w); // cm == 0 when j == 0
      Wt[li] = w;
      T::dot_acc(lj, xc[li], w);
      T::dot_acc(lm, vm, w);
    } else {
      w = Wt[li];
    }
#pragma unroll
    for (int i = 0; i < NC; ++i) T::dot_acc(la + i * W, v[i], w);
  }
#pragma unroll
  for (int k = 0; k < NA * W; ++k) acc[k] = la[k];
#pragma unroll
  for (int k = 0; k < W; ++k) {
    accM[k] = lm[k];
    accA[k] = lj[k];
  }
}

template <class S, bool ANISO, int NC, bool FIRST>
__device__ __forceinline__ void pk_phase_a(const PkArgs &a, PkCtx<S> &c, GridSync &gs, S *start, int j, int i0,
                                           double inv_bj, double cm) {
  using T = Sc<S>;
  constexpr int W = T::W;
  constexpr int NA = NC > 0 ? NC : 1;
  double acc[NA * W], accM[W], accA[W];
  pk_loop_a<S, ANISO, NC, FIRST>(a.g, c.cnt, j, i0, c.X() + a.g.nx, c.Wt(), c.flag(), start + c.s0,
                                 (const S *)a.V + c.s0, a.colstride, ANISO ? a.cf + c.s0 : nullptr, inv_bj, cm,
                                 acc, accM, accA);
  PKT(900 + j); // phase A site loop done
  if (FIRST && j > 0 && NC == j - 1) {
    // the usual case (all other columns in this group): partial-sum slots 0..j are contiguous -> one publish
    double all[(NC + 2) * W];
#pragma unroll
    for (int k = 0; k < NC * W; ++k) all[k] = acc[k];
#pragma unroll
    for (int k = 0; k < W; ++k) {
      all[NC * W + k] = accM[k];
      all[(NC + 1) * W + k] = accA[k];
    }
    pk_publish<(NC + 2) * W>(all, a.part, gs.slot, c.red(), 0);
    return;
  }
  if (NC > 0) pk_publish<NA * W>(acc, a.part, gs.slot, c.red(), i0 * W);
  if (FIRST) {
    if (j > 0) pk_publish<W>(accM, a.part, gs.slot, c.red(), (j - 1) * W);
    pk_publish<W>(accA, a.part, gs.slot, c.red(), j * W);
  }
}

// Phase B, group of NC other columns.  LAST: also subtracts the j-1 and j terms, accumulates
// ||w||^2 and stores w to column j+1 (global) and to the X tile centre.
template <class S, int NC, bool LAST>
__device__ __forceinline__ double pk_loop_b(const int cnt, const int j, const int i0, S *__restrict__ xc,
                                            S *__restrict__ Wt, const S *__restrict__ startp,
                                            const S *__restrict__ vbase, const size_t cs, S *__restrict__ coln,
                                            const S *__restrict__ gco) {
#ifdef PK_HINT_B
  PK_LSH(xc); PK_LSH(Wt); PK_LSH(gco); PK_LGL(startp); PK_LGL(vbase); PK_LGL(coln);
#endif
  using T = Sc<S>;
  constexpr int W = T::W;
  constexpr int NA = NC > 0 ? NC : 1;
  const S *p[NA];
  S gc[NA];
#pragma unroll
  for (int i = 0; i < NC; ++i) {
    p[i] = pk_colp<S>(startp, vbase, cs, i0 + i);
    gc[i] = gco[i0 + i];
  }
  const S *colm = j > 0 ? pk_colp<S>(startp, vbase, cs, j - 1) : startp;
  const S gm = j > 0 ? gco[j - 1] : T::zero(), gj = gco[j];
  double nrm = 0.0;
#pragma unroll(pk_unroll(NC, W))
  for (int li = threadIdx.x; li < cnt; li += kPkThreads) {
    S v[NA];
#pragma unroll
    for (int i = 0; i < NC; ++i) v[i] = __ldcg(p[i] + li);
    S vm = T::zero();
    if (LAST && j > 0) vm = __ldcg(colm + li);
    S w = Wt[li];
#pragma unroll
    for (int i = 0; i < NC; ++i) w = T::sub(w, T::mul(gc[i], v[i]));
    if (LAST) {
      w = T::sub(w, T::mul(gm, vm));
      w = T::sub(w, T::mul(gj, xc[li]));
      nrm += T::abs2(w);
      coln[li] = w;
      xc[li] = w;
    } else {
      Wt[li] = w;
    }
  }
  return nrm;
}

template <class S, int NC, bool LAST>
__device__ __forceinline__ void pk_phase_b(const PkArgs &a, PkCtx<S> &c, GridSync &gs, S *start, int j, int i0) {
  double nrm[1];
  nrm[0] = pk_loop_b<S, NC, LAST>(c.cnt, j, i0, c.X() + a.g.nx, c.Wt(), start + c.s0, (const S *)a.V + c.s0,
                                  a.colstride, (S *)a.V + (size_t)j * a.colstride + c.s0, c.gco());
  PKT(800 + j); // phase B site loop done
  if (LAST) pk_publish<1>(nrm, a.part, gs.slot, c.red(), 0);
}

#define PK_SWITCH_NC(nc, CALL)                                                                     \
  switch (nc) {                                                                                    \
  case 0: { constexpr int NC = 0; CALL; } break;                                                   \
  case 1: { constexpr int NC = 1; CALL; } break;                                                   \
  case 2: { constexpr int NC = 2; CALL; } break;                                                   \
  case 3: { constexpr int NC = 3; CALL; } break;                                                   \
  case 4: { constexpr int NC = 4; CALL; } break;                                                   \
  case 5: { constexpr int NC = 5; CALL; } break;                                                   \
  case 6: { constexpr int NC = 6; CALL; } break;                                                   \
  case 7: { constexpr int NC = 7; CALL; } break;                                                   \
  default: { constexpr int NC = 8; CALL; } break;                                                  \
  }

// ---------------------------------------------------------------- one Lanczos run
// On entry: X centre holds the start vector's owned sites, nrm2[0] = ||start||^2 (gathered),
// the start vector is complete in global memory (a barrier has passed since it was written).
// On exit: columns 0..m-1 complete in global memory, alpha[0..m-2], nrm2[0..m-1] in shared.
template <class S, bool ANISO>
__device__ bool pk_lanczos(const PkArgs &a, PkCtx<S> &c, GridSync &gs, S *start) {
  using T = Sc<S>;
  constexpr int W = T::W;
  for (int j = 0; j < a.m - 1; ++j) {
    if (j == 0) pk_load_halo<S>(a, c, pk_col<S>(a, start, 0)); // later iterations: loaded with the norm gather
    PKT(100 + j); // halo loaded
    const double bj = sqrt(c.nrm2()[j]);
    const double inv_bj = 1.0 / bj;
    const double cm = j > 0 ? -bj / sqrt(c.nrm2()[j - 1]) : 0.0; // -beta_j * (1/beta_{j-1})
    const int nother = j > 0 ? j - 1 : 0;
    // ---- phase A: w = L v_j - beta_j v_{j-1};  sums_i = w_i^H w for i <= j
    {
      const int nc0 = min(nother, kPkNV);
      PK_SWITCH_NC(nc0, (pk_phase_a<S, ANISO, NC, true>(a, c, gs, start, j, 0, inv_bj, cm)));
      for (int i0 = kPkNV; i0 < nother; i0 += kPkNV) {
        const int nc = min(nother - i0, kPkNV);
        PK_SWITCH_NC(nc, (pk_phase_a<S, ANISO, NC, false>(a, c, gs, start, j, i0, inv_bj, cm)));
      }
    }
    PKT(200 + j); // phase A done (loop + publish)
    if (!grid_barrier(gs)) return false;
    PKT(300 + j); // barrier passed
    pk_gather((j + 1) * W, a.part, gs.slot ^ 1, a.ncta, c.vals());
    // h_i = sums_i / beta_i ; w -= h_i v_i = (sums_i / beta_i^2) w_i ; alpha = Re h_j
    if (threadIdx.x <= j) {
      const int i = threadIdx.x;
      c.gco()[i] = T::mulr(T::from(c.vals() + i * W), 1.0 / c.nrm2()[i]);
      if (i == j) c.alpha()[j] = c.vals()[j * W] * inv_bj; // T(j,j) = Re(v_j^H w)
    }
    __syncthreads();
    PKT(400 + j); // gathered
    // ---- phase B: w -= sum_i g_i w_i ; ||w||^2 ; w -> column j+1 (global) and X centre
    {
      int i0 = 0;
      for (; nother - i0 > kPkNV; i0 += kPkNV) pk_phase_b<S, kPkNV, false>(a, c, gs, start, j, i0);
      PK_SWITCH_NC(nother - i0, (pk_phase_b<S, NC, true>(a, c, gs, start, j, i0)));
    }
    PKT(500 + j); // phase B done
    if (!grid_barrier(gs)) return false;
    PKT(600 + j); // barrier passed
    // one latency window: warp 0 gathers ||w||^2 while warps 1.. fetch the next iteration's halo rows
    if (j + 1 < a.m - 1 && threadIdx.x >= 32)
      pk_load_halo_part<S>(a, c, pk_col<S>(a, start, j + 1), threadIdx.x - 32, kPkThreads - 32);
    else
      pk_gather_nosync(1, a.part, gs.slot ^ 1, a.ncta, c.vals());
    if (threadIdx.x == 0) c.nrm2()[j + 1] = c.vals()[0];
    __syncthreads();
    PKT(700 + j); // gathered
  }
  return true;
}

// warp 0: eigen-decomposition of T and the coefficient vectors
//   co_f[i] = beta_0 * (Q f(t Lambda) Q^T e_1)_i / beta_i      (columns are un-normalised)
template <class S>
__device__ __forceinline__ void pk_coefficients(const PkArgs &a, PkCtx<S> &c, int nfn, int f0, double t0r,
                                                double t0i, int f1, double t1r, double t1i) {
  const int m = a.m;
#ifndef PK_QL_REG
#define PK_QL_REG 1
#endif
  if (PK_QL_REG && threadIdx.x < 32 && m <= 32) {
    // register-resident QL (nls_kernels.cuh): lane i owns d_i, e_i; ~3x shorter than the shared-memory form
    const int lane = threadIdx.x;
    PkTri &T = c.tri;
    double *const tz = T.Z;
    const int ldz = T.ldz;
    double dr = lane < m - 1 ? c.alpha()[lane] : 0.0; // T(m-1,m-1) = 0
    double er = lane < m - 1 ? sqrt(c.nrm2()[lane + 1]) : 0.0;
    if (lane < m)
      for (int k = 0; k < m; ++k) tz[lane * ldz + k] = lane == k ? 1.0 : 0.0;
    __syncwarp();
    const bool ok = tridiag_warp_reg(dr, er, tz, ldz, m);
    __syncwarp();
    if (!ok && lane == 0) atomicExch(a.status, 1);
    TridiagArgs ta;
    ta.m = m; ta.W = 1; ta.nfn = nfn;
    ta.fn[0] = f0; ta.tr[0] = t0r; ta.ti[0] = t0i;
    ta.fn[1] = f1; ta.tr[1] = t1r; ta.ti[1] = t1i;
    const double scale = lane < m ? sqrt(c.nrm2()[0]) / sqrt(c.nrm2()[lane]) : 0.0;
    tridiag_coefs_reg(dr, tz, ldz, m, ta, scale, T.e, c.coef(), 2 * kMaxM);
  } else if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    PkTri &T = c.tri;
    double *const td = T.d, *const te = T.e, *const tz = T.Z;
    PK_SHARED(td);
    PK_SHARED(te);
    PK_SHARED(tz);
    const int ldz = T.ldz;
    for (int i = 0; i < m; ++i) {
      td[i * 32 + lane] = i < m - 1 ? c.alpha()[i] : 0.0; // T(m-1,m-1) = 0
      te[i * 32 + lane] = i < m - 1 ? sqrt(c.nrm2()[i + 1]) : 0.0;
    }
    for (int i = lane; i < m; i += 32)
      for (int k = 0; k < m; ++k) tz[i * ldz + k] = i == k ? 1.0 : 0.0;
    __syncwarp();
    const bool ok = pk_tridiag_warp(T, m);
    __syncwarp();
    if (!ok && lane == 0) atomicExch(a.status, 1);
    const double beta0 = sqrt(c.nrm2()[0]);
    for (int f = 0; f < nfn; ++f) {
      const int fn = f ? f1 : f0;
      const double tr = f ? t1r : t0r, ti = f ? t1i : t0i;
      double *co = c.coef() + f * 2 * kMaxM;
      __syncwarp();
      for (int k = lane; k < m; k += 32) { // f(t lambda_k) -> e[k][0..1] (e is free after QL)
        const double2 fv = fn_eval(fn, tr, ti, td[k * 32 + lane]);
        te[k * 32] = fv.x;
        te[k * 32 + 1] = fv.y;
      }
      __syncwarp();
      for (int i = lane; i < m; i += 32) {
        double sr = 0.0, si = 0.0;
        for (int k = 0; k < m; ++k) {
          const double zz = tz[i * ldz + k] * tz[k];
          sr = fma(zz, te[k * 32], sr);
          si = fma(zz, te[k * 32 + 1], si);
        }
        const double scale = beta0 / sqrt(c.nrm2()[i]);
        co[2 * i] = sr * scale;
        co[2 * i + 1] = si * scale;
      }
    }
  }
  __syncthreads();
}

// ---------------------------------------------------------------- recombination
// y += sum_{q<NC} co[q] * base[q*cs]   (base already points at the site)
template <int NC>
__device__ __forceinline__ void pk_comb(const double *base, size_t cs, const double *co, double &y) {
  double v[NC];
#pragma unroll
  for (int q = 0; q < NC; ++q) v[q] = __ldcg(base + q * cs);
#pragma unroll
  for (int q = 0; q < NC; ++q) y = fma(v[q], co[2 * q], y);
}
template <int NC>
__device__ __forceinline__ void pk_comb(const double2 *base, size_t cs, const double *co, double2 &y) {
  double2 v[NC];
#pragma unroll
  for (int q = 0; q < NC; ++q) v[q] = __ldcg(base + q * cs);
#pragma unroll
  for (int q = 0; q < NC; ++q) {
    const double cr = co[2 * q], ci = co[2 * q + 1];
    y.x = fma(v[q].x, cr, fma(-v[q].y, ci, y.x));
    y.y = fma(v[q].x, ci, fma(v[q].y, cr, y.y));
  }
}
template <int NC>
__device__ __forceinline__ void pk_comb2(const double *base, size_t cs, const double *co0, const double *co1,
                                         double &y0, double &y1) {
  double v[NC];
#pragma unroll
  for (int q = 0; q < NC; ++q) v[q] = __ldcg(base + q * cs);
#pragma unroll
  for (int q = 0; q < NC; ++q) {
    y0 = fma(v[q], co0[2 * q], y0);
    y1 = fma(v[q], co1[2 * q], y1);
  }
}

// y(off) = sum_i co_i * column_i[off]: column 0 = start buffer, columns 1..m-1 strided in V
template <class S>
__device__ __forceinline__ S pk_combine(const PkArgs &a, const S *start, const double *co, long long off) {
  S y = Sc<S>::zero();
  pk_comb<1>(start + off, 0, co, y);
  const S *vb = (const S *)a.V + off;
  const int nv = a.m - 1;
  int q = 0;
  for (; q + 4 <= nv; q += 4) pk_comb<4>(vb + (size_t)q * a.colstride, a.colstride, co + 2 * (1 + q), y);
  switch (nv - q) {
  case 1: pk_comb<1>(vb + (size_t)q * a.colstride, a.colstride, co + 2 * (1 + q), y); break;
  case 2: pk_comb<2>(vb + (size_t)q * a.colstride, a.colstride, co + 2 * (1 + q), y); break;
  case 3: pk_comb<3>(vb + (size_t)q * a.colstride, a.colstride, co + 2 * (1 + q), y); break;
  default: break;
  }
  return y;
}
__device__ __forceinline__ void pk_combine2(const PkArgs &a, const double *start, const double *co0,
                                            const double *co1, long long off, double &y0, double &y1) {
  y0 = 0.0;
  y1 = 0.0;
  pk_comb2<1>(start + off, 0, co0, co1, y0, y1);
  const double *vb = (const double *)a.V + off;
  const int nv = a.m - 1;
  int q = 0;
  for (; q + 4 <= nv; q += 4)
    pk_comb2<4>(vb + (size_t)q * a.colstride, a.colstride, co0 + 2 * (1 + q), co1 + 2 * (1 + q), y0, y1);
  switch (nv - q) {
  case 1: pk_comb2<1>(vb + (size_t)q * a.colstride, a.colstride, co0 + 2 * (1 + q), co1 + 2 * (1 + q), y0, y1); break;
  case 2: pk_comb2<2>(vb + (size_t)q * a.colstride, a.colstride, co0 + 2 * (1 + q), co1 + 2 * (1 + q), y0, y1); break;
  case 3: pk_comb2<3>(vb + (size_t)q * a.colstride, a.colstride, co0 + 2 * (1 + q), co1 + 2 * (1 + q), y0, y1); break;
  default: break;
  }
}

// ---------------------------------------------------------------- kernel set-up
inline size_t pk_smem_bytes(size_t esz, long long csz, int nx, int m, int wt_halo = 0) {
  size_t b = (size_t)(2 * csz + 2 * nx + wt_halo) * esz; // X + Wt (+ halo rows of the Wt tile, Stoermer-Verlet)
  b += ((size_t)csz + 15) / 16 * 16;                      // flags
  b += (size_t)m * 32 * 8 * 2 + (size_t)m * (m + 1) * 8;  // tridiagonal work space
  b += (size_t)(2 * kPkNV + 4) * kPkWarps * 8;            // red
  b += (size_t)kPkMaxK * 8 + kMaxM * 8 + (kMaxM + 2) * 8 + 4 * kMaxM * 8 + kMaxM * 16;
  return b + 64;
}

template <class S>
__device__ __forceinline__ void pk_carve(const PkArgs &a, PkCtx<S> &c, unsigned char *smem, int wt_halo = 0) {
  const long long tile = a.csz + 2 * a.g.nx;
  c.X_ = (S *)smem;
  c.Wt_ = c.X_ + tile;
  c.flag_ = (unsigned char *)(c.Wt_ + a.csz + wt_halo);
  double *p = (double *)(c.flag() + (a.csz + 15) / 16 * 16);
  c.tri.d = p; p += a.m * 32;
  c.tri.e = p; p += a.m * 32;
  c.tri.Z = p; p += a.m * (a.m + 1);
  c.tri.ldz = a.m + 1;
  c.red_ = (double (*)[kPkWarps])p; p += (2 * kPkNV + 4) * kPkWarps;
  c.vals_ = p; p += kPkMaxK;
  c.alpha_ = p; p += kMaxM;
  c.nrm2_ = p; p += kMaxM + 2; // even count: keeps the complex gco array 16-byte aligned without integer casts
  c.coef_ = p; p += 4 * kMaxM;
  c.gco_ = (S *)p;
  c.s0 = (long long)blockIdx.x * a.csz;
  long long e0 = c.s0 + a.csz;
  if (e0 > a.g.n) e0 = a.g.n;
  c.cnt = e0 > c.s0 ? (int)(e0 - c.s0) : 0;
  for (int li = threadIdx.x; li < c.cnt; li += kPkThreads) c.flag()[li] = pk_site_flags(a.g, c.s0 + li);
  __syncthreads();
}

// publish ||start||^2, barrier, gather -> nrm2[0]
template <class S>
__device__ __forceinline__ bool pk_start_norm(const PkArgs &a, PkCtx<S> &c, GridSync &gs, double (&nrm)[1]) {
  pk_publish<1>(nrm, a.part, gs.slot, c.red(), 0);
  if (!grid_barrier(gs)) return false;
  pk_gather(1, a.part, gs.slot ^ 1, a.ncta, c.vals());
  if (threadIdx.x == 0) c.nrm2()[0] = c.vals()[0];
  __syncthreads();
  return true;
}

__device__ __forceinline__ double2 pk_nlse_phase(int eq, const NlseParams &p, double mv, double2 dens, double2 z) {
  switch (eq) {
  case NLS_EQ_NLSE_CUBIC: return nlse_phase<NLS_EQ_NLSE_CUBIC>(p, mv, dens, z, z);
  case NLS_EQ_NLSE_CQ: return nlse_phase<NLS_EQ_NLSE_CQ>(p, mv, dens, z, z);
  default: return nlse_phase<NLS_EQ_NLSE_SAT>(p, mv, dens, z, z);
  }
}
__device__ __forceinline__ double pk_f_real(int eq, double x) {
  switch (eq) {
  case NLS_EQ_SGE: return f_real<NLS_EQ_SGE>(x);
  case NLS_EQ_KGE: return f_real<NLS_EQ_KGE>(x);
  case NLS_EQ_PHI4: return f_real<NLS_EQ_PHI4>(x);
  case NLS_EQ_SGE_DOUBLE: return f_real<NLS_EQ_SGE_DOUBLE>(x);
  default: return f_real<NLS_EQ_SGE_SINH>(x);
  }
}

// ---------------------------------------------------------------- NLSE, Strang splitting
// buffers: u ping-pong (a.u0 current at entry, a.u1), start-vector ping-pong (sv0, sv1)
template <bool ANISO>
__global__ void __launch_bounds__(kPkThreads, 1) k_pk_nlse(PkArgs a, double2 *sv0, double2 *sv1) {
  extern __shared__ __align__(16) unsigned char pk_smem[];
  PkCtx<double2> c;
  pk_carve<double2>(a, c, pk_smem);
  GridSync gs{a.bar, 0u, (unsigned)a.ncta, 0, a.status};
  if (a.debug_fault && blockIdx.x == (unsigned)a.ncta - 1u) return; // test hook: one CTA never arrives
  const NlseParams p{0.5 * a.dt, a.s1, a.s2, a.kappa};
  double2 *ucur = (double2 *)a.u0, *unext = (double2 *)a.u1;
  double2 *scur = sv0, *snext = sv1;
  PK_GLOBAL(ucur);
  PK_GLOBAL(unext);
  PK_GLOBAL(scur);
  PK_GLOBAL(snext);
  double2 *xc = c.X() + a.g.nx;
  const double *mfc = a.mf + c.s0;
  // start vector of the first step: rho = exp(.5 tau g(u)) u   (nlse_cubic_solver.hpp:62-66)
  {
    double nrm[1] = {0.0};
    for (int li = threadIdx.x; li < c.cnt; li += kPkThreads) {
      const double2 uv = ucur[c.s0 + li];
      const double2 r = pk_nlse_phase(a.eq, p, mfc[li], uv, uv);
      scur[c.s0 + li] = r;
      xc[li] = r;
      nrm[0] += r.x * r.x + r.y * r.y;
    }
    if (!pk_start_norm<double2>(a, c, gs, nrm)) return;
  }
  for (int step = 0; step < a.nsteps; ++step) {
    if (!pk_lanczos<double2, ANISO>(a, c, gs, scur)) return;
    pk_coefficients<double2>(a, c, 1, NLS_FN_EXP, 0.0, a.dt, 0, 0.0, 0.0); // expm_multiply(L, rho, i dt)
    double nrm[1] = {0.0};
    for (int li = threadIdx.x; li < c.cnt; li += kPkThreads) {
      const long long idx = c.s0 + li;
      // fused neumann_bc_no_velocity (boundaries.hpp:40-57): boundary sites take the clamped site's value
      const long long sc = (c.flag()[li] & PF_RING) ? clamp_site(a.g, idx) : idx;
      const double2 buf = pk_combine<double2>(a, scur, c.coef(), sc);
      const double mv = a.mf[sc];
      double2 un;
      if (a.eq == NLS_EQ_NLSE_CUBIC) un = nlse_phase<NLS_EQ_NLSE_CUBIC>(p, mv, buf, buf, buf);
      else un = pk_nlse_phase(a.eq, p, mv, __ldcg(ucur + sc), buf); // old-u density
      unext[idx] = un;
      const double2 r = pk_nlse_phase(a.eq, p, mfc[li], un, un);
      snext[idx] = r;
      xc[li] = r;
      nrm[0] += r.x * r.x + r.y * r.y;
    }
    if (!pk_start_norm<double2>(a, c, gs, nrm)) return;
    double2 *t = ucur; ucur = unext; unext = t;
    t = scur; scur = snext; snext = t;
  }
}

// Site loops of the Gautschi step (restrict parameters, see pk_loop_a).
// mid: two coefficient sets in one pass over the basis; g = -m f(filter u) -> gbuf + X tile, ||g||^2,
//      acc = 2 cos(.)u - u_past, u_past <- u                      (sg_single_solver.hpp:46-52)
__device__ __forceinline__ double pk_gautschi_mid(const PkArgs &a, const int cnt, const long long s0,
                                                  const double *__restrict__ u, const double *__restrict__ V,
                                                  const double *__restrict__ coef, const double *__restrict__ mfc,
                                                  double *__restrict__ gbuf, double *__restrict__ xc,
                                                  double *__restrict__ accb, double *__restrict__ up) {
#ifdef PK_HINT_G
  PK_LSH(coef); PK_LSH(xc); PK_LGL(u); PK_LGL(V); PK_LGL(mfc); PK_LGL(gbuf); PK_LGL(accb); PK_LGL(up);
#endif
  const int m = a.m, eq = a.eq;
  const size_t cs = a.colstride;
  double co0[2 * kPkCoefRegs], co1[2 * kPkCoefRegs]; // unused beyond m; m <= kPkCoefRegs uses registers
  const bool inreg = m <= kPkCoefRegs;
#pragma unroll
  for (int q = 0; q < kPkCoefRegs; ++q) {
    co0[q] = q < m ? coef[2 * q] : 0.0;
    co1[q] = q < m ? coef[2 * kMaxM + 2 * q] : 0.0;
  }
  double nrm = 0.0;
#pragma unroll(kPkCombUnroll)
  for (int li = threadIdx.x; li < cnt; li += kPkThreads) {
    const long long idx = s0 + li;
    double sf = 0.0, sc = 0.0;
    if (inreg) {
      double v[kPkCoefRegs];
      v[0] = __ldcg(u + idx);
#pragma unroll
      for (int q = 1; q < kPkCoefRegs; ++q) v[q] = q < m ? __ldcg(V + (size_t)(q - 1) * cs + idx) : 0.0;
#pragma unroll
      for (int q = 0; q < kPkCoefRegs; ++q) {
        sf = fma(v[q], co0[q], sf);
        sc = fma(v[q], co1[q], sc);
      }
    } else {
      pk_combine2(a, u, coef, coef + 2 * kMaxM, idx, sf, sc);
    }
    const double gv = -(mfc[li] * pk_f_real(eq, sf));
    gbuf[idx] = gv;
    xc[li] = gv;
    nrm += gv * gv;
    accb[idx] = 2 * sc - __ldcg(up + li);
    up[li] = __ldcg(u + idx);
  }
  return nrm;
}
// tail: u+ = acc + tau^2 sinc^2(.) g at the clamped site (fused no-flux copy), ||u+||^2   (:53-57, boundaries.hpp:40-57)
__device__ __forceinline__ double pk_gautschi_tail(const PkArgs &a, const int cnt, const long long s0,
                                                   const double *__restrict__ gbuf, const double *__restrict__ V,
                                                   const double *__restrict__ coef,
                                                   const unsigned char *__restrict__ flag,
                                                   const double *__restrict__ accb, double *__restrict__ u,
                                                   double *__restrict__ xc, const double tau2) {
#ifdef PK_HINT_G
  PK_LSH(coef); PK_LSH(xc); PK_LSH(flag); PK_LGL(u); PK_LGL(V); PK_LGL(gbuf); PK_LGL(accb);
#endif
  const int m = a.m;
  const size_t cs = a.colstride;
  double co[kPkCoefRegs];
  const bool inreg = m <= kPkCoefRegs;
#pragma unroll
  for (int q = 0; q < kPkCoefRegs; ++q) co[q] = q < m ? coef[2 * q] : 0.0;
  double nrm = 0.0;
#pragma unroll(kPkCombUnroll)
  for (int li = threadIdx.x; li < cnt; li += kPkThreads) {
    const long long idx = s0 + li;
    const long long sc = (flag[li] & PF_RING) ? clamp_site(a.g, idx) : idx;
    double s = 0.0;
    if (inreg) {
      double v[kPkCoefRegs];
      v[0] = __ldcg(gbuf + sc);
#pragma unroll
      for (int q = 1; q < kPkCoefRegs; ++q) v[q] = q < m ? __ldcg(V + (size_t)(q - 1) * cs + sc) : 0.0;
#pragma unroll
      for (int q = 0; q < kPkCoefRegs; ++q) s = fma(v[q], co[q], s);
    } else {
      s = pk_combine<double>(a, gbuf, coef, sc);
    }
    const double un = __ldcg(accb + sc) + tau2 * s;
    u[idx] = un;
    xc[li] = un;
    nrm += un * un;
  }
  return nrm;
}

// ---------------------------------------------------------------- real equations, Gautschi
// buffers: a.u0 = u (also column 0 of the first Lanczos run), a.u1 = u_past, a.acc, gbuf = start
// vector of the second run.   sg_single_solver.hpp:43-59, kg_solver.hpp:13-22
template <bool ANISO>
__global__ void __launch_bounds__(kPkThreads, 1) k_pk_gautschi(PkArgs a, double *gbuf) {
  extern __shared__ __align__(16) unsigned char pk_smem[];
  PkCtx<double> c;
  pk_carve<double>(a, c, pk_smem);
  GridSync gs{a.bar, 0u, (unsigned)a.ncta, 0, a.status};
  if (a.debug_fault && blockIdx.x == (unsigned)a.ncta - 1u) return; // test hook: one CTA never arrives
  double *u = (double *)a.u0, *up = (double *)a.u1 + c.s0, *accb = (double *)a.acc;
  PK_GLOBAL(u);
  PK_GLOBAL(up);
  PK_GLOBAL(accb);
  PK_GLOBAL(gbuf);
  double *xc = c.X() + a.g.nx;
  const double *mfc = a.mf + c.s0;
  const int filt = a.eq == NLS_EQ_SGE ? NLS_FN_MOD_COSINE : NLS_FN_ID_SQRT;
  const double tau2 = a.dt * a.dt;
  {
    double nrm[1] = {0.0};
    for (int li = threadIdx.x; li < c.cnt; li += kPkThreads) {
      const double v = u[c.s0 + li];
      xc[li] = v;
      nrm[0] += v * v;
    }
    if (!pk_start_norm<double>(a, c, gs, nrm)) return;
  }
  for (int step = 0; step < a.nsteps; ++step) {
    if (step == a.nsteps - 1) PKT_INIT();
    PKT(1);
    if (!pk_lanczos<double, ANISO>(a, c, gs, u)) return;
    PKT(2); // first Lanczos run done
    pk_coefficients<double>(a, c, 2, filt, a.dt, 0.0, NLS_FN_COS_SQRT, a.dt, 0.0);
    PKT(3); // tridiagonal solve + coefficients
    {
      double nrm[1];
      nrm[0] = pk_gautschi_mid(a, c.cnt, c.s0, u, (const double *)a.V, c.coef(), mfc, gbuf, xc, accb, up);
      if (!pk_start_norm<double>(a, c, gs, nrm)) return;
    }
    PKT(4); // combine 1 + start norm
    if (!pk_lanczos<double, ANISO>(a, c, gs, gbuf)) return;
    PKT(5);
    pk_coefficients<double>(a, c, 1, NLS_FN_SINC2_HALF, a.dt, 0.0, 0, 0.0, 0.0);
    PKT(6);
    {
      double nrm[1];
      nrm[0] = pk_gautschi_tail(a, c.cnt, c.s0, gbuf, (const double *)a.V, c.coef(), c.flag(), accb, u, xc, tau2);
      if (!pk_start_norm<double>(a, c, gs, nrm)) return;
    }
    PKT(7); // combine 2 + start norm
  }
  PKT_DUMP(a.ncta);
}


// ---------------------------------------------------------------- real equations, Stoermer-Verlet
// u+ = 2u - u_past + tau^2 (L u - m f(u)) followed by the no-flux copy (sg_single_sv_solver.hpp:9-19,
// kg_sv_solver.hpp:13-21, boundaries.hpp:40-57), nsteps steps per launch.  u (with its +-nx halo
// rows) and u_past stay in shared memory for the whole launch; each step writes u+ once to the
// rotating global buffers (the neighbours' halo rows and the final state come from there) and
// costs one grid barrier.  buf[0] = u, buf[2] = u_past at entry; step s writes buf[(s+1)%3].
// one Stoermer-Verlet step over a chunk (restrict parameters, see pk_loop_a): reads the u tile (with halo
// rows) and u_past in place, writes u+ over u_past in shared memory and to the rotating global buffer
template <bool ANISO>
__device__ __forceinline__ void pk_sv_loop(const PkArgs &a, const int cnt, const long long s0,
                                           const double *__restrict__ xc, double *__restrict__ pc,
                                           const unsigned char *__restrict__ flag,
                                           const double *__restrict__ ucur, const double *__restrict__ upast,
                                           double *__restrict__ unew, const double *__restrict__ mf,
                                           const double *__restrict__ cfg, const double tau2) {
  PK_LSH(xc); PK_LSH(pc); PK_LSH(flag); PK_LGL(unew);
  const int eq = a.eq;
#pragma unroll 2
  for (int li = threadIdx.x; li < cnt; li += kPkThreads) {
    const long long idx = s0 + li;
    const unsigned f = flag[li];
    double un;
    if (f & PF_RING) { // boundary site: the value the update gives at the clamped (interior) site
      const long long sc = clamp_site(a.g, idx);
      const double uc = __ldcg(ucur + sc);
      const double lu = stencil_site_f<double, ANISO>(a.g, [&](long long q) { return __ldcg(ucur + q); }, a.cf, sc, 1.0);
      un = 2 * uc - __ldcg(upast + sc) + tau2 * (lu - mf[sc] * pk_f_real(eq, uc));
    } else {
      const double uc = xc[li];
      const double lu = pk_stencil<double, ANISO>(a.g, xc, ucur + s0, cfg, f, li);
      un = 2 * uc - pc[li] + tau2 * (lu - mf[idx] * pk_f_real(eq, uc));
    }
    pc[li] = un;
    unew[idx] = un;
  }
}

// halo rows of `col` into the tile whose centre is xc (tile = xc - nx)
__device__ __forceinline__ void pk_sv_halo(const PkArgs &a, const int cnt, const long long s0, double *xc,
                                           const double *col) {
  const int nx = a.g.nx;
  constexpr int kQ = 4;
  if (cnt > 0) {
    for (int t0 = threadIdx.x; t0 < 2 * nx; t0 += kQ * kPkThreads) {
      double v[kQ];
      int xi[kQ];
#pragma unroll
      for (int q = 0; q < kQ; ++q) {
        const int t = t0 + q * kPkThreads;
        const bool up = t >= nx;
        const long long idx = up ? s0 + cnt + (t - nx) : s0 - nx + t;
        xi[q] = INT_MIN;
        if (t < 2 * nx && idx >= 0 && idx < a.g.n) {
          xi[q] = up ? cnt + (t - nx) : t - nx;
          v[q] = __ldcg(col + idx);
        }
      }
#pragma unroll
      for (int q = 0; q < kQ; ++q)
        if (xi[q] != INT_MIN) xc[xi[q]] = v[q];
    }
  }
  __syncthreads();
}

template <bool ANISO>
__global__ void __launch_bounds__(kPkThreads, 1) k_pk_sv(PkArgs a, double *b0, double *b1, double *b2) {
  extern __shared__ __align__(16) unsigned char pk_smem[];
  PkCtx<double> c;
  pk_carve<double>(a, c, pk_smem, 2 * a.g.nx); // both tiles carry halo rows: they swap roles every step
  GridSync gs{a.bar, 0u, (unsigned)a.ncta, 0, a.status};
  if (a.debug_fault && blockIdx.x == (unsigned)a.ncta - 1u) return; // test hook: one CTA never arrives
  double *buf[3] = {b0, b1, b2};
  double *xc = c.X() + a.g.nx, *pc = c.Wt() + a.g.nx; // u tile centre, u_past tile centre
  const double *cfg = ANISO ? a.cf + c.s0 : nullptr;
  const double tau2 = a.dt * a.dt;
  const int cnt = c.cnt;
  const long long s0 = c.s0;
  for (int li = threadIdx.x; li < cnt; li += kPkThreads) {
    xc[li] = buf[0][s0 + li];
    pc[li] = buf[2][s0 + li];
  }
  __syncthreads();
  int p = 0;
  for (int step = 0; step < a.nsteps; ++step) {
    const double *ucur = buf[p];
    pk_sv_halo(a, cnt, s0, xc, ucur);
    pk_sv_loop<ANISO>(a, cnt, s0, xc, pc, c.flag(), ucur, buf[(p + 2) % 3], buf[(p + 1) % 3], a.mf, cfg, tau2);
    double *t = xc; xc = pc; pc = t; // u+ (written over u_past) becomes u, u becomes u_past
    if (!grid_barrier(gs)) return;   // starts with __syncthreads: all tile reads of this step are done
    p = (p + 1) % 3;
  }
}


// ---------------------------------------------------------------- Stoermer-Verlet, 2D, temporally blocked ("sv2")
// u+ = 2u - u_past + tau^2 (L u - m f(u)) followed by the no-flux copy (sg_single_sv_solver.hpp:9-19,
// kg_sv_solver.hpp:13-21, boundaries.hpp:40-57) for 2D grids with the uniform Laplacian, nsteps steps per launch.
// One CTA per SM owns a w x h TILE of the grid; u, u_past and m of the tile PLUS a halo ring of depth s stay in
// shared memory for the whole launch.  The CTA advances s time steps between two exchanges: step t (1..s)
// updates the tile grown by s - t rings (redundant rim compute, the valid region shrinks by one ring per step),
// so per s steps there is ONE neighbour synchronisation and ONE halo exchange through L2 -- no HBM traffic at
// all in steady state, no grid-wide barrier: a CTA only waits for the progress counters of its <= 8 neighbour
// tiles (release/acquire).  Tiles are written to two global buffer sets in turn (set b % 2 after block b), so a
// neighbour still reading block b's halo never races with block b+1's stores.
// Only true interior sites are ever evaluated (boundary sites take the value of their clamped interior site, the
// fused form of neumann_bc_no_velocity), so the stencil needs none of the boundary-row logic: all four neighbours
// exist and the diagonal is -4 (laplacians.hpp:23-30 applies -3 to ring rows only).
struct Sv2Plan {
  int px, py, w, h, s; // tiles in x / y, tile extent, temporal block depth
  size_t smem;
};
struct Sv2Args {
  int nx, ny, nsteps;
  Sv2Plan pl;
  double sc, tau2;
  const double *mf;
  double *su[3], *sp[3];     // buffer sets (u, u_past): [0] = input, [1], [2] = ping-pong outputs
  unsigned long long *prog;  // [ncta] blocks completed by each CTA (zeroed before the launch)
  unsigned *abort_flag;
  int *status;
  int debug_fault;
};

inline size_t sv2_smem_bytes(int w, int h, int s) { return (size_t)3 * (w + 2 * s) * (h + 2 * s) * 8 + 64; }

// host: choose the tiling and the block depth.  Cost model in site updates per step and SM: the grown regions
// average (w+s)(h+s) sites, one exchange (neighbour wait + halo reload + tile store) costs about kSync updates.
inline bool sv2_plan(int nx, int ny, int nsm, size_t smem_max, Sv2Plan &best) {
  const double kSync = 6000.0;
  double best_cost = 1e300;
  bool found = false;
  for (int py = 1; py <= nsm; ++py) {
    int px = nsm / py;
    if (px < 1) break;
    int w = (nx + px - 1) / px, h = (ny + py - 1) / py;
    if (w < 4) w = nx < 4 ? nx : 4;
    if (h < 4) h = ny < 4 ? ny : 4;
    px = (nx + w - 1) / w;
    const int pyy = (ny + h - 1) / h;
    if ((long long)px * pyy > nsm) continue;
    // the last tile of a row / column may be narrower: it must still hold the clamped site of its boundary sites
    if ((nx - (px - 1) * w) < 2 && px > 1) continue;
    if ((ny - (pyy - 1) * h) < 2 && pyy > 1) continue;
    const int lastw = nx - (px - 1) * w, lasth = ny - (pyy - 1) * h;
    for (int s = 1; s <= 8; ++s) {
      if (s > (lastw < w ? lastw : w) || s > (lasth < h ? lasth : h)) break; // halos reach adjacent tiles only
      const size_t sm = sv2_smem_bytes(w, h, s);
      if (sm > smem_max) break;
      const double cost = (double)(w + s) * (h + s) + kSync / s;
      if (cost < best_cost) {
        best_cost = cost;
        best = Sv2Plan{px, pyy, w, h, s, sm};
        found = true;
      }
    }
  }
  return found;
}

template <int EQ>
__global__ void __launch_bounds__(kPkThreads, 1) k_pk_sv2(Sv2Args a) {
  extern __shared__ __align__(16) unsigned char pk_smem[];
  __shared__ int s_ok;
  const int w = a.pl.w, h = a.pl.h, S = a.pl.s, nx = a.nx, ny = a.ny;
  const int P = w + 2 * S, H = h + 2 * S;
  double *A = (double *)pk_smem, *B = A + (size_t)P * H, *M = B + (size_t)P * H;
  const int tx = blockIdx.x % a.pl.px, ty = blockIdx.x / a.pl.px;
  const int x0 = tx * w, x1 = min(nx, x0 + w), y0 = ty * h, y1 = min(ny, y0 + h);
  const int lx = threadIdx.x & 31, ly = threadIdx.x >> 5;
  const int org = S * P + S - y0 * P - x0; // smem index of grid site (x, y) = org + y*P + x
  if (a.debug_fault && blockIdx.x == gridDim.x - 1) return; // test hook: this CTA never reports progress
  // whole grown tile of u, u_past (input set) and m
  {
    const int gx0 = max(0, x0 - S), gx1 = min(nx, x1 + S), gy0 = max(0, y0 - S), gy1 = min(ny, y1 + S);
    for (int y = gy0 + ly; y < gy1; y += kPkWarps)
      for (int x = gx0 + lx; x < gx1; x += 32) {
        const size_t gi = (size_t)y * nx + x;
        const int si = org + y * P + x;
        A[si] = __ldcg(a.su[0] + gi);
        B[si] = __ldcg(a.sp[0] + gi);
        M[si] = __ldcg(a.mf + gi);
      }
  }
  __syncthreads();
  const double sc = a.sc, tau2 = a.tau2, dg = -4.0 * a.sc;
  int done = 0;
  for (int b = 0; done < a.nsteps; ++b) {
    const int sb = min(S, a.nsteps - done);
    if (b > 0) {
      // wait for the <= 8 neighbour tiles to have stored block b-1, then refresh the halo rings from their set
      if (threadIdx.x < 9 && threadIdx.x != 4) {
        const int nxt = tx + (int)(threadIdx.x % 3) - 1, nyt = ty + (int)(threadIdx.x / 3) - 1;
        if (nxt >= 0 && nxt < a.pl.px && nyt >= 0 && nyt < a.pl.py) {
          const unsigned long long *pp = a.prog + (nyt * a.pl.px + nxt);
          const long long t0 = clock64();
          while (ld_acquire_u64(pp) < (unsigned long long)b) {
            if (clock64() - t0 > 4000000000LL || ld_acquire_u32(a.abort_flag) != 0u) { // ~2 s
              atomicExch(a.abort_flag, 1u);
              atomicExch(a.status, 4); // poisons the handle (nls.h)
              break;
            }
          }
        }
      }
      __syncthreads();
      const double *ru = a.su[1 + ((b - 1) & 1)], *rp = a.sp[1 + ((b - 1) & 1)];
      const int gx0 = max(0, x0 - sb), gx1 = min(nx, x1 + sb), gy0 = max(0, y0 - sb), gy1 = min(ny, y1 + sb);
      for (int y = gy0 + ly; y < gy1; y += kPkWarps) {
        const bool yin = y >= y0 && y < y1;
        for (int x = gx0 + lx; x < gx1; x += 32) {
          if (yin && x >= x0 && x < x1) continue; // own tile: already in shared memory
          const size_t gi = (size_t)y * nx + x;
          const int si = org + y * P + x;
          A[si] = __ldcg(ru + gi);
          B[si] = __ldcg(rp + gi);
        }
      }
      if (threadIdx.x == 0) s_ok = ld_acquire_u32(a.abort_flag) == 0u;
      __syncthreads();
      if (!s_ok) return; // uniform: the whole CTA unwinds
    }
    for (int t = 1; t <= sb; ++t) {
      const int g = sb - t;
      const int rx0 = max(0, x0 - g), rx1 = min(nx, x1 + g), ry0 = max(0, y0 - g), ry1 = min(ny, y1 + g);
      // pass 1: interior sites of the domain inside the region
      const int ix0 = max(rx0, 1), ix1 = min(rx1, nx - 1), iy0 = max(ry0, 1), iy1 = min(ry1, ny - 1);
      for (int y = iy0 + ly; y < iy1; y += kPkWarps) {
        const int rowi = org + y * P;
#pragma unroll 2
        for (int x = ix0 + lx; x < ix1; x += 32) {
          const int si = rowi + x;
          const double uc = A[si];
          double acc = sc * A[si - P];           // same accumulation order as stencil_site_f: y-, x-, centre, x+, y+
          acc = fma(sc, A[si - 1], acc);
          acc = fma(dg, uc, acc);
          acc = fma(sc, A[si + 1], acc);
          acc = fma(sc, A[si + P], acc);
          const double buf = acc - M[si] * f_real<EQ>(uc);
          B[si] = 2 * uc - B[si] + tau2 * buf;   // u+ over u_past, in place
        }
      }
      __syncthreads();
      // pass 2: boundary sites of the domain inside the region take their clamped site's new value
      if (rx0 == 0 || rx1 == nx || ry0 == 0 || ry1 == ny) {
        if (rx0 == 0 || rx1 == nx)
          for (int y = ry0 + (int)threadIdx.x; y < ry1; y += kPkThreads) {
            const int yc = min(max(y, 1), ny - 2);
            if (rx0 == 0) B[org + y * P + 0] = B[org + yc * P + 1];
            if (rx1 == nx) B[org + y * P + nx - 1] = B[org + yc * P + nx - 2];
          }
        if (ry0 == 0 || ry1 == ny)
          for (int x = rx0 + (int)threadIdx.x; x < rx1; x += kPkThreads) {
            const int xc = min(max(x, 1), nx - 2);
            if (ry0 == 0) B[org + 0 * P + x] = B[org + 1 * P + xc];
            if (ry1 == ny) B[org + (ny - 1) * P + x] = B[org + (ny - 2) * P + xc];
          }
        __syncthreads();
      }
      double *tmp = A; A = B; B = tmp; // u+ becomes u, u becomes u_past
    }
    // store the tile (u and u_past) to this block's set; neighbours read their halo rings from it
    {
      double *wu = a.su[1 + (b & 1)], *wp = a.sp[1 + (b & 1)];
      for (int y = y0 + ly; y < y1; y += kPkWarps)
        for (int x = x0 + lx; x < x1; x += 32) {
          const size_t gi = (size_t)y * nx + x;
          const int si = org + y * P + x;
          wu[gi] = A[si];
          wp[gi] = B[si];
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) st_release_u64(a.prog + blockIdx.x, (unsigned long long)(b + 1));
    done += sb;
  }
}

} // namespace nls
