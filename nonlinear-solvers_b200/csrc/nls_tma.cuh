// nls_tma.cuh -- TMA-staged forms of the two Lanczos passes (single trajectory, large grids).
//
// The passes of nls_kernels.cuh keep every column value of a site in registers, so the number
// of loads a thread can have in flight is bounded by its register budget (complex, 12 columns:
// 48 registers of data + 48 of accumulators => 2 CTAs/SM and latency-bound stencil loads).
// Here the data movement is taken off the threads: one elected thread per CTA issues 1-D bulk
// tensor-memory-accelerator copies (cp.async.bulk.shared.global, completion counted on an
// mbarrier) that stream TILE-site segments of w and of every basis column into a ring of
// shared-memory stages, several tiles ahead of the arithmetic.  The arithmetic then runs out of
// shared memory in two access patterns that registers could not serve at once:
//   * site-parallel   (thread = site):  w -= sum_i h_i c_i, the stencil, norms
//   * column-parallel (warp = column):  s_i = c_i^H w, g_i = c_i^H w''  (2 accumulators/thread)
// Same arithmetic per site as k_lanczos_a / k_lanczos_b; sums are formed in a different order.
#pragma once
#include "nls_kernels.cuh"

namespace nls {

constexpr int kTmaThreads = 256;
constexpr int kTmaWarps = kTmaThreads / 32;
constexpr int kTmaMaxCols = 12;
constexpr int kTmaTileBytes = 4096; // bytes per column segment: 256 complex / 512 real sites

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try(unsigned long long *bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a byte-count mismatch must never hang the GPU.  After ~2 s the CTA gives up, raises
// the handle's status word (reported by nls_sync) and skips every later wait.
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity, volatile int *s_abort, int *status) {
  if (*s_abort) return;
  if (mbar_try(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {
      *s_abort = 1;
      atomicExch(status, 3);
      return;
    }
  }
}
// 1-D bulk copy global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src, unsigned bytes, unsigned long long *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---------------------------------------------------------------- CTA totals -> grid totals
// Like grid_reduce, but the CTA's K totals are already in shared memory (s_cta[K]).
template <int KMAX>
__device__ __forceinline__ void cta_totals_reduce(int K, const double *s_cta, const RedCtx &rc, double *dst,
                                                  double *dst1, int split) {
  __shared__ double s_tot[KMAX];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y, nblk = gridDim.x;
  double *mine = rc.partials + ((size_t)b * nblk + blockIdx.x) * K;
  if ((int)threadIdx.x < K) mine[threadIdx.x] = s_cta[threadIdx.x];
  if (rc.slab != nullptr) __threadfence_system();
  else __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned t = atomicAdd(&rc.tickets[b], 1u);
    s_last = (t == (unsigned)nblk - 1u);
    if (s_last) rc.tickets[b] = 0u;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const double *all = rc.partials + (size_t)b * nblk * K;
  const bool slab = rc.slab != nullptr;
  for (int k = warp; k < K; k += kTmaWarps) {
    double v = 0.0;
    for (int blk = lane; blk < nblk; blk += 32) v += __ldcg(&all[(size_t)blk * K + k]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) {
      if (slab) s_tot[k] = v;
      else if (k < split) dst[k] = v;
      else dst1[k - split] = v;
    }
  }
  if (slab) {
    __syncthreads();
    slab_allreduce_rt(*rc.slab, K, s_tot, dst, dst1, split);
  }
}

struct TmaPlan {
  int tile;     // sites per tile
  int nstages;  // ring depth
  int ntile;    // tiles per trajectory
  size_t stage_bytes, smem_bytes;
};


// shared-memory layout (dynamic): [mbar: 8 x 8 B][h: 12 S][s_cta: 32 doubles][pad to 128]
//                                 [stage 0: (ncol + nextra) segments of kTmaTileBytes (+ small tails)] ...
constexpr int kTmaHdrBytes = 1024;

// ---------------------------------------------------------------- pass B
// w -= sum_i h_i column_i over the columns [i0, i0+cnt); LAST: ||w||^2 -> nrm2[j+1], the Gram
// entries g_i = column_i^H w of the finished column, T(j,j), halo push.   (k_lanczos_b)
template <class S, bool LAST>
__global__ void __launch_bounds__(kTmaThreads, 1)
k_lanczos_b_tma(Geom g, TmaPlan pl, const S *__restrict__ start, S *__restrict__ V, size_t colstride, int j,
                int i0, int cnt, double *scal, RedCtx rc, Halo<S> hp, int *status) {
  using T = Sc<S>;
  extern __shared__ __align__(128) unsigned char tma_smem[];
  unsigned long long *bar = (unsigned long long *)tma_smem;
  volatile int *s_abort = (volatile int *)(tma_smem + 56);
  S *h = (S *)(tma_smem + 64);
  double *s_cta = (double *)(tma_smem + 64 + kTmaMaxCols * 16);
  unsigned char *stages = tma_smem + kTmaHdrBytes;
  double *sb = scal + (size_t)blockIdx.y * kScalStride;
  S *wc = traj(V + (size_t)j * colstride, g);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int TL = pl.tile;

  if (tid < kTmaMaxCols) { // same coefficients as k_lanczos_b
    S hv = T::zero();
    if (tid < cnt) {
      const int ci = i0 + tid;
      S si = T::from(sb + kOffH + ci * T::W);
      if (ci < j) {
        const double q = sb[kOffH + j * T::W] / sb[kOffNrm2 + j];
        si = T::sub(si, T::mulr(T::from(sb + kOffG + ci * T::W), q));
      }
      hv = T::mulr(si, 1.0 / sb[kOffNrm2 + ci]);
    }
    h[tid] = hv;
  }
  if (tid == 0) {
    *s_abort = 0;
    for (int s = 0; s < pl.nstages; ++s) mbar_init(bar + s, 1);
    fence_proxy_async();
  }
  __syncthreads();

  // producer: segment 0 = w (column j+1), segments 1..cnt = columns i0..i0+cnt-1
  auto issue = [&](int it) {
    const long long tile = (long long)blockIdx.x + (long long)it * gridDim.x;
    if (tile >= pl.ntile) return;
    const int s = it % pl.nstages;
    const long long s0 = tile * TL;
    const int len = (int)min((long long)TL, g.n - s0);
    const unsigned bytes = (unsigned)(len * sizeof(S));
    unsigned char *st = stages + (size_t)s * pl.stage_bytes;
    mbar_expect_tx(bar + s, bytes * (unsigned)(cnt + 1));
    bulk_g2s(st, wc + s0, bytes, bar + s);
    for (int i = 0; i < cnt; ++i)
      bulk_g2s(st + (size_t)(i + 1) * kTmaTileBytes, traj(colptr<S>(start, V, colstride, i0 + i), g) + s0, bytes, bar + s);
  };
  if (tid == 0)
    for (int it = 0; it < pl.nstages; ++it) issue(it);

  double nrm = 0.0;
  double ga[2 * T::W]; // Gram partials of the (at most two) columns this warp owns
#pragma unroll
  for (int k = 0; k < 2 * T::W; ++k) ga[k] = 0.0;
  for (int it = 0;; ++it) {
    const long long tile = (long long)blockIdx.x + (long long)it * gridDim.x;
    if (tile >= pl.ntile) break;
    const int s = it % pl.nstages;
    const long long s0 = tile * TL;
    const int len = (int)min((long long)TL, g.n - s0);
    S *wt = (S *)(stages + (size_t)s * pl.stage_bytes);
    mbar_wait(bar + s, (unsigned)((it / pl.nstages) & 1), s_abort, status);
    for (int t = tid; t < len; t += kTmaThreads) { // site-parallel update
      S wv = wt[t];
#pragma unroll 4
      for (int i = 0; i < cnt; ++i) wv = T::sub(wv, T::mul(h[i], ((const S *)((unsigned char *)wt + (size_t)(i + 1) * kTmaTileBytes))[t]));
      wc[s0 + t] = wv;
      if (LAST) {
        nrm += T::abs2(wv);
        halo_push<S>(hp, g.n, g.plane, s0 + t, wv);
        wt[t] = wv;
      }
    }
    if (LAST) {
      __syncthreads();
#pragma unroll
      for (int c = 0; c < 2; ++c) { // column-parallel Gram entries
        const int i = warp + c * kTmaWarps;
        if (i < cnt) {
          const S *ct = (const S *)((unsigned char *)wt + (size_t)(i + 1) * kTmaTileBytes);
          for (int t = lane; t < len; t += 32) T::dot_acc(ga + c * T::W, ct[t], wt[t]);
        }
      }
    }
    __syncthreads(); // every thread is done with this stage
    if (tid == 0) {
      fence_proxy_async();
      issue(it + pl.nstages);
    }
  }
  if (!LAST) return;
  // CTA totals: s_cta[0 .. cnt*W) = g, s_cta[cnt*W] = ||w||^2
  __shared__ double s_w[kTmaWarps];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) nrm += __shfl_down_sync(0xffffffffu, nrm, o);
  if (lane == 0) s_w[warp] = nrm;
#pragma unroll
  for (int c = 0; c < 2; ++c)
#pragma unroll
    for (int k = 0; k < T::W; ++k) {
      double v = ga[c * T::W + k];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
      const int i = warp + c * kTmaWarps;
      if (lane == 0 && i < cnt) s_cta[i * T::W + k] = v;
    }
  __syncthreads();
  if (tid == 0) {
    double v = 0.0;
    for (int w = 0; w < kTmaWarps; ++w) v += s_w[w];
    s_cta[cnt * T::W] = v;
    if (blockIdx.x == 0) sb[kOffAlpha + j * T::W] = sb[kOffH + j * T::W] / sqrt(sb[kOffNrm2 + j]);
  }
  __syncthreads();
  cta_totals_reduce<kTmaMaxCols * 2 + 1>(cnt * T::W + 1, s_cta, rc, sb + kOffG + i0 * T::W, sb + kOffNrm2 + j + 1,
                                         cnt * T::W);
}


// ---------------------------------------------------------------- pass A
// w = L v_j - beta_j v_{j-1} into column j+1 and s_i = column_i^H w for ALL columns 0..j (cnt = j+1
// <= kTmaMaxCols; full re-orthogonalisation).  Stage segments (kTmaTileBytes apart):
//   [0 .. cnt)   columns 0..j          (column j doubles as the stencil centre, j-1 as the recurrence term)
//   cnt+0..+3    column j at -nx, +nx, -plane, +plane
//   cnt+4        w (written by the stencil phase, read by the dot phase)
//   cnt+5..+9    ANISO: c at 0, -nx, +nx, -plane, +plane   (8-byte elements)
//   tail         4 x 16 B: the x-neighbours just outside the tile (column j left/right, c left/right)
template <class S, bool ANISO>
__global__ void __launch_bounds__(kTmaThreads, 1)
k_lanczos_a_tma(Geom g, TmaPlan pl, const S *__restrict__ start, S *__restrict__ V, size_t colstride, int j,
                const double *__restrict__ cf, double *scal, RedCtx rc, int *status) {
  using T = Sc<S>;
  constexpr int EPS = 16 / (int)sizeof(S); // elements per 16 bytes (1 complex, 2 real)
  extern __shared__ __align__(128) unsigned char tma_smem[];
  unsigned long long *bar = (unsigned long long *)tma_smem;
  volatile int *s_abort = (volatile int *)(tma_smem + 56);
  double *s_cta = (double *)(tma_smem + 64 + kTmaMaxCols * 16);
  unsigned char *stages = tma_smem + kTmaHdrBytes;
  double *sb = scal + (size_t)blockIdx.y * kScalStride;
  const int cnt = j + 1;
  const S *cj = traj(colptr<S>(start, V, colstride, j), g);
  S *wc = traj(V + (size_t)j * colstride, g);
  const double *cfb = ANISO ? cf + g.halo : nullptr;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int TL = pl.tile;
  const double bj = sqrt(sb[kOffNrm2 + j]);
  const double inv = 1.0 / bj;
  const double cmul = j > 0 ? -bj / sqrt(sb[kOffNrm2 + j - 1]) : 0.0;
  const long long lo_lim = -g.halo, hi_lim = g.n + g.halo; // addressable range of a vector (ghost planes incl.)
  const size_t tail_off = (size_t)(cnt + (ANISO ? 10 : 5)) * kTmaTileBytes;

  if (tid == 0) {
    *s_abort = 0;
    for (int s = 0; s < pl.nstages; ++s) mbar_init(bar + s, 1);
    fence_proxy_async();
  }
  __syncthreads();

  auto issue = [&](int it) {
    const long long tile = (long long)blockIdx.x + (long long)it * gridDim.x;
    if (tile >= pl.ntile) return;
    const int s = it % pl.nstages;
    const long long s0 = tile * TL;
    const int len = (int)min((long long)TL, g.n - s0);
    unsigned char *st = stages + (size_t)s * pl.stage_bytes;
    unsigned total = 0;
    // clipped segment copy: elements [a, a+len) of p (element size esz) -> dst (+ the clipped-off prefix)
    auto seg = [&](unsigned char *dst, const void *p, int esz, long long a, bool count_only) {
      long long lo = a < lo_lim ? lo_lim : a, hi = a + len > hi_lim ? hi_lim : a + len;
      if (hi <= lo) return;
      const unsigned bytes = (unsigned)((hi - lo) * esz);
      if (count_only) total += bytes;
      else bulk_g2s(dst + (size_t)(lo - a) * esz, (const char *)p + lo * esz, bytes, bar + s);
    };
    for (int pass = 0; pass < 2; ++pass) { // pass 0: count bytes and arm the barrier; pass 1: issue
      const bool co = pass == 0;
      for (int i = 0; i < cnt; ++i)
        seg(st + (size_t)i * kTmaTileBytes, traj(colptr<S>(start, V, colstride, i), g), sizeof(S), s0, co);
      seg(st + (size_t)(cnt + 0) * kTmaTileBytes, cj, sizeof(S), s0 - g.nx, co);
      seg(st + (size_t)(cnt + 1) * kTmaTileBytes, cj, sizeof(S), s0 + g.nx, co);
      if (g.dim3) {
        seg(st + (size_t)(cnt + 2) * kTmaTileBytes, cj, sizeof(S), s0 - g.plane, co);
        seg(st + (size_t)(cnt + 3) * kTmaTileBytes, cj, sizeof(S), s0 + g.plane, co);
      }
      if (ANISO) {
        seg(st + (size_t)(cnt + 5) * kTmaTileBytes, cfb, 8, s0, co);
        seg(st + (size_t)(cnt + 6) * kTmaTileBytes, cfb, 8, s0 - g.nx, co);
        seg(st + (size_t)(cnt + 7) * kTmaTileBytes, cfb, 8, s0 + g.nx, co);
        if (g.dim3) {
          seg(st + (size_t)(cnt + 8) * kTmaTileBytes, cfb, 8, s0 - g.plane, co);
          seg(st + (size_t)(cnt + 9) * kTmaTileBytes, cfb, 8, s0 + g.plane, co);
        }
      }
      // 16-byte edge pieces: [s0-EPS, s0) and [s0+len, s0+len+EPS) of column j, [s0-2, s0) / [s0+len, +2) of c
      if (s0 - EPS >= lo_lim) { if (co) total += 16; else bulk_g2s(st + tail_off, cj + (s0 - EPS), 16, bar + s); }
      if (s0 + len + EPS <= hi_lim) { if (co) total += 16; else bulk_g2s(st + tail_off + 16, cj + (s0 + len), 16, bar + s); }
      if (ANISO) {
        if (s0 - 2 >= lo_lim) { if (co) total += 16; else bulk_g2s(st + tail_off + 32, cfb + (s0 - 2), 16, bar + s); }
        if (s0 + len + 2 <= hi_lim) { if (co) total += 16; else bulk_g2s(st + tail_off + 48, cfb + (s0 + len), 16, bar + s); }
      }
      if (co) mbar_expect_tx(bar + s, total);
    }
  };
  if (tid == 0)
    for (int it = 0; it < pl.nstages; ++it) issue(it);

  double sa[2 * T::W];
#pragma unroll
  for (int k = 0; k < 2 * T::W; ++k) sa[k] = 0.0;
  for (int it = 0;; ++it) {
    const long long tile = (long long)blockIdx.x + (long long)it * gridDim.x;
    if (tile >= pl.ntile) break;
    const int s = it % pl.nstages;
    const long long s0 = tile * TL;
    const int len = (int)min((long long)TL, g.n - s0);
    unsigned char *st = stages + (size_t)s * pl.stage_bytes;
    const S *xc = (const S *)(st + (size_t)j * kTmaTileBytes);
    const S *xm1 = (const S *)(st + (size_t)(j > 0 ? j - 1 : 0) * kTmaTileBytes);
    const S *xym = (const S *)(st + (size_t)(cnt + 0) * kTmaTileBytes), *xyp = (const S *)(st + (size_t)(cnt + 1) * kTmaTileBytes);
    const S *xzm = (const S *)(st + (size_t)(cnt + 2) * kTmaTileBytes), *xzp = (const S *)(st + (size_t)(cnt + 3) * kTmaTileBytes);
    S *wt = (S *)(st + (size_t)(cnt + 4) * kTmaTileBytes);
    const double *cc = (const double *)(st + (size_t)(cnt + 5) * kTmaTileBytes);
    const double *cym = (const double *)(st + (size_t)(cnt + 6) * kTmaTileBytes), *cyp = (const double *)(st + (size_t)(cnt + 7) * kTmaTileBytes);
    const double *czm = (const double *)(st + (size_t)(cnt + 8) * kTmaTileBytes), *czp = (const double *)(st + (size_t)(cnt + 9) * kTmaTileBytes);
    const S *edge = (const S *)(st + tail_off);            // [0..EPS): left piece, [EPS..2 EPS): right piece
    const double *cedge = (const double *)(st + tail_off + 32); // [0,1]: c[s0-2], c[s0-1]; [2,3]: c[s0+len], c[s0+len+1]
    mbar_wait(bar + s, (unsigned)((it / pl.nstages) & 1), s_abort, status);
    for (int t = tid; t < len; t += kTmaThreads) { // site-parallel: stencil + recurrence term
      const unsigned f = site_flags(g, s0 + t);
      const S z = T::zero();
      const S vxm = (f & NB_XM) ? (t > 0 ? xc[t - 1] : edge[EPS - 1]) : z;
      const S vxp = (f & NB_XP) ? (t < len - 1 ? xc[t + 1] : edge[EPS]) : z;
      const S vym = (f & NB_YM) ? xym[t] : z, vyp = (f & NB_YP) ? xyp[t] : z;
      const S vzm = (f & NB_ZM) ? xzm[t] : z, vzp = (f & NB_ZP) ? xzp[t] : z;
      double c0 = 0, kxm = 0, kxp = 0, kym = 0, kyp = 0, kzm = 0, kzp = 0;
      if (ANISO) {
        c0 = cc[t];
        kxm = (f & NB_XM) ? (t > 0 ? cc[t - 1] : cedge[1]) : 0.0;
        kxp = (f & NB_XP) ? (t < len - 1 ? cc[t + 1] : cedge[2]) : 0.0;
        kym = (f & NB_YM) ? cym[t] : 0.0;
        kyp = (f & NB_YP) ? cyp[t] : 0.0;
        kzm = (f & NB_ZM) ? czm[t] : 0.0;
        kzp = (f & NB_ZP) ? czp[t] : 0.0;
      }
      S w = stencil_vals<S, ANISO>(g, f, inv, xc[t], vzm, vym, vxm, vxp, vyp, vzp, c0, kzm, kym, kxm, kxp, kyp, kzp);
      if (j > 0) w = T::fma_r(cmul, xm1[t], w);
      wc[s0 + t] = w;
      wt[t] = w;
    }
    __syncthreads();
#pragma unroll
    for (int c = 0; c < 2; ++c) { // column-parallel: s_i = column_i^H w
      const int i = warp + c * kTmaWarps;
      if (i < cnt) {
        const S *ct = (const S *)(st + (size_t)i * kTmaTileBytes);
        for (int t = lane; t < len; t += 32) T::dot_acc(sa + c * T::W, ct[t], wt[t]);
      }
    }
    __syncthreads();
    if (tid == 0) {
      fence_proxy_async();
      issue(it + pl.nstages);
    }
  }
#pragma unroll
  for (int c = 0; c < 2; ++c)
#pragma unroll
    for (int k = 0; k < T::W; ++k) {
      double v = sa[c * T::W + k];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
      const int i = warp + c * kTmaWarps;
      if (lane == 0 && i < cnt) s_cta[i * T::W + k] = v;
    }
  __syncthreads();
  cta_totals_reduce<kTmaMaxCols * 2 + 1>(cnt * T::W, s_cta, rc, sb + kOffH, sb + kOffH, cnt * T::W);
}

} // namespace nls
