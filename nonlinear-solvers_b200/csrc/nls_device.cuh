// nls_device.cuh -- device-side building blocks: scalar traits, grid geometry, the
// matrix-free no-flux stencil (K1) and the deterministic grid-wide reduction.
//
// The stencil reproduces, row for row, the sparse matrices the reference assembles in
// nlsolvers/common/include/laplacians.hpp:9-218 (uniform and anisotropic, 2D and 3D),
// including its quirks: it is written on the FLAT index,
// exactly like the reference's triplet loops, so the corner rows (-3 diagonal with two
// neighbours) and the 3D "y-neighbour without plane check" coupling fall out of the same
// conditions (idx+nx < N, (idx+1) % nx != 0, ...) instead of being special-cased.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace nls {

constexpr int kThreads = 256;
constexpr int kMaxM = 64;
constexpr int kMaxBlocksPerTraj = 148 * 4; // persistent-style grid: 4 CTAs of 256 thr per SM

// ---------------------------------------------------------------- scalar traits
template <class S> struct Sc;
template <> struct Sc<double> {
  static constexpr int W = 1; // doubles per scalar
  __device__ static double zero() { return 0.0; }
  __device__ static double mulr(double a, double r) { return a * r; }
  __device__ static double add(double a, double b) { return a + b; }
  __device__ static double sub(double a, double b) { return a - b; }
  __device__ static double mul(double a, double b) { return a * b; }
  __device__ static double fma_r(double r, double a, double acc) { return fma(r, a, acc); } // acc + r*a
  __device__ static double abs2(double a) { return a * a; }
  // acc[0..W) += conj(a) * b
  __device__ static void dot_acc(double *acc, double a, double b) { acc[0] = fma(a, b, acc[0]); }
  __device__ static double from(const double *p) { return p[0]; }
  __device__ static double conj(double a) { return a; }
  __device__ static double real(double r) { return r; }
  __device__ static double re(double a) { return a; }
};
template <> struct Sc<double2> {
  static constexpr int W = 2;
  __device__ static double2 zero() { return make_double2(0.0, 0.0); }
  __device__ static double2 mulr(double2 a, double r) { return make_double2(a.x * r, a.y * r); }
  __device__ static double2 add(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
  __device__ static double2 sub(double2 a, double2 b) { return make_double2(a.x - b.x, a.y - b.y); }
  __device__ static double2 mul(double2 a, double2 b) {
    return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
  }
  __device__ static double2 fma_r(double r, double2 a, double2 acc) {
    return make_double2(fma(r, a.x, acc.x), fma(r, a.y, acc.y));
  }
  __device__ static double abs2(double2 a) { return a.x * a.x + a.y * a.y; }
  __device__ static void dot_acc(double *acc, double2 a, double2 b) {
    acc[0] = fma(a.x, b.x, fma(a.y, b.y, acc[0]));
    acc[1] = fma(a.x, b.y, fma(-a.y, b.x, acc[1]));
  }
  __device__ static double2 from(const double *p) { return make_double2(p[0], p[1]); }
  __device__ static double2 conj(double2 a) { return make_double2(a.x, -a.y); }
  __device__ static double2 real(double r) { return make_double2(r, 0.0); }
  __device__ static double re(double2 a) { return a.x; }
};

// ---------------------------------------------------------------- geometry
struct Geom {
  long long n;     // sites owned by this handle (per trajectory)
  long long nglob; // sites of the global grid (== n unless z-slab decomposed)
  long long goff;  // global flat index of local site 0 (z0 * nx * ny)
  long long plane; // nx*ny (0 in 2D)
  int nx, ny, nz_loc, nz_glob;
  int z0;
  int dim3;      // 0: 2D, 1: 3D
  long long halo; // ghost elements in front of (and behind) every trajectory's array
  long long stride; // elements between consecutive trajectories (= n + 2*halo)
  double sc;     // op_sign / (dx*dy)   [2D]   or   op_sign / (dx*dx)   [3D]
  int pf;        // grid sweeps the flat Lanczos passes prefetch ahead into L2 (0 = off; pf_l2 below)
  int pf_first;  // the same for the stencil-carrying flat pass A
  int pf_fin;    // the same for the recombination kernel (k_nlse_finish)
  int raw;       // pass A writes the RAW vector L column_j (no 1/beta, no recurrence term): NLS_REORTH_GRAM
};

// The flat passes are one batch of loads per thread and grid sweep: a sweep lasts one loaded DRAM round trip
// (~2 us at 16 warps/SM, profiles/r02j_ncu_full_c3_pass_a_tm_v1_vs_flat.txt: long-scoreboard 11 per issue, issue
// slots 27 % busy).  Prefetching the lines of the thread's NEXT sweep into L2 costs no registers and turns the
// exposed latency into an L2 hit.  Lead of one sweep by default (nls_engine.cu: kDefaultPrefetch, measurements there).
__device__ __forceinline__ void pf_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// y = (L x)(idx) for one site.  x points at local site 0 (ghost planes, if any, sit at
// negative indices / beyond n).  cf (ANISO only) is laid out like x.
// x is any callable long long -> S (global pointer, or the shared-memory tile of the persistent
// kernel); it is only evaluated at neighbours that exist.
template <class S, bool ANISO, class X>
__device__ __forceinline__ S stencil_site_f(const Geom &g, X x, const double *__restrict__ cf,
                                            long long idx, double pre) {
  using T = Sc<S>;
  const long long gi = idx + g.goff;
  const unsigned ugi = (unsigned)gi;
  const unsigned i = ugi % (unsigned)g.nx;
  const bool has_xm = i != 0;                      // (gi) % nx != 0       laplacians.hpp:33
  const bool has_xp = i != (unsigned)(g.nx - 1);   // (gi+1) % nx != 0
  const bool has_ym = gi >= g.nx;                  // loop i < N-nx, no plane check (:39, :140)
  const bool has_yp = gi + g.nx < g.nglob;
  const bool has_zm = g.dim3 && gi >= g.plane;     // :145
  const bool has_zp = g.dim3 && gi + g.plane < g.nglob;
  const S xc = T::mulr(x(idx), pre);
  S acc = T::zero();
  if (!ANISO) {
    double dg;
    if (!g.dim3) {
      const unsigned j = ugi / (unsigned)g.nx;
      const bool ring = i == 0 || i == (unsigned)(g.nx - 1) || j == 0 || j == (unsigned)(g.ny - 1);
      dg = ring ? -3.0 : -4.0; // :23-30
    } else {
      const unsigned jj = ugi / (unsigned)g.nx;
      const unsigned j = jj % (unsigned)g.ny, k = jj / (unsigned)g.ny;
      const bool face = i == 0 || i == (unsigned)(g.nx - 1) || j == 0 || j == (unsigned)(g.ny - 1) ||
                        k == 0 || k == (unsigned)(g.nz_glob - 1);
      dg = face ? -5.0 : -6.0; // :124-127
    }
    // accumulate in increasing column order, like the CSC product of the reference
    if (has_zm) acc = T::fma_r(g.sc, T::mulr(x(idx - g.plane), pre), acc);
    if (has_ym) acc = T::fma_r(g.sc, T::mulr(x(idx - g.nx), pre), acc);
    if (has_xm) acc = T::fma_r(g.sc, T::mulr(x(idx - 1), pre), acc);
    acc = T::fma_r(dg * g.sc, xc, acc);
    if (has_xp) acc = T::fma_r(g.sc, T::mulr(x(idx + 1), pre), acc);
    if (has_yp) acc = T::fma_r(g.sc, T::mulr(x(idx + g.nx), pre), acc);
    if (has_zp) acc = T::fma_r(g.sc, T::mulr(x(idx + g.plane), pre), acc);
  } else {
    const double cc = cf[idx];
    // weights (c_a + c_b)/2 and the diagonal sum in the reference's accumulation order:
    // x-links (i-1,i),(i,i+1); y-links; z-links   (:75-95, :181-210)
    const double wxm = has_xm ? (cf[idx - 1] + cc) / 2.0 : 0.0;
    const double wxp = has_xp ? (cc + cf[idx + 1]) / 2.0 : 0.0;
    const double wym = has_ym ? (cf[idx - g.nx] + cc) / 2.0 : 0.0;
    const double wyp = has_yp ? (cc + cf[idx + g.nx]) / 2.0 : 0.0;
    const double wzm = has_zm ? (cf[idx - g.plane] + cc) / 2.0 : 0.0;
    const double wzp = has_zp ? (cc + cf[idx + g.plane]) / 2.0 : 0.0;
    double ds = 0.0;
    if (has_xm) ds += wxm;
    if (has_xp) ds += wxp;
    if (has_ym) ds += wym;
    if (has_yp) ds += wyp;
    if (has_zm) ds += wzm;
    if (has_zp) ds += wzp;
    if (has_zm) acc = T::fma_r(wzm * g.sc, T::mulr(x(idx - g.plane), pre), acc);
    if (has_ym) acc = T::fma_r(wym * g.sc, T::mulr(x(idx - g.nx), pre), acc);
    if (has_xm) acc = T::fma_r(wxm * g.sc, T::mulr(x(idx - 1), pre), acc);
    acc = T::fma_r(-ds * g.sc, xc, acc);
    if (has_xp) acc = T::fma_r(wxp * g.sc, T::mulr(x(idx + 1), pre), acc);
    if (has_yp) acc = T::fma_r(wyp * g.sc, T::mulr(x(idx + g.nx), pre), acc);
    if (has_zp) acc = T::fma_r(wzp * g.sc, T::mulr(x(idx + g.plane), pre), acc);
  }
  return acc;
}

// The same rows from neighbour VALUES (missing neighbours passed as anything, gated by the flags):
// used by the plane-marching kernels, which keep the z+-1 values in registers and fetch the
// in-plane neighbours through L1.  Accumulation order identical to stencil_site_f.
enum { NB_XM = 1, NB_XP = 2, NB_YM = 4, NB_YP = 8, NB_ZM = 16, NB_ZP = 32, NB_RING = 64 };
template <class S, bool ANISO>
__device__ __forceinline__ S stencil_vals(const Geom &g, unsigned f, double pre, S xc, S xzm, S xym, S xxm,
                                          S xxp, S xyp, S xzp, double cc, double czm, double cym,
                                          double cxm, double cxp, double cyp, double czp) {
  using T = Sc<S>;
  S acc = T::zero();
  xc = T::mulr(xc, pre);
  if (!ANISO) {
    const double dg = g.dim3 ? ((f & NB_RING) ? -5.0 : -6.0) : ((f & NB_RING) ? -3.0 : -4.0);
    if (f & NB_ZM) acc = T::fma_r(g.sc, T::mulr(xzm, pre), acc);
    if (f & NB_YM) acc = T::fma_r(g.sc, T::mulr(xym, pre), acc);
    if (f & NB_XM) acc = T::fma_r(g.sc, T::mulr(xxm, pre), acc);
    acc = T::fma_r(dg * g.sc, xc, acc);
    if (f & NB_XP) acc = T::fma_r(g.sc, T::mulr(xxp, pre), acc);
    if (f & NB_YP) acc = T::fma_r(g.sc, T::mulr(xyp, pre), acc);
    if (f & NB_ZP) acc = T::fma_r(g.sc, T::mulr(xzp, pre), acc);
  } else {
    const double wxm = (f & NB_XM) ? (cxm + cc) / 2.0 : 0.0;
    const double wxp = (f & NB_XP) ? (cc + cxp) / 2.0 : 0.0;
    const double wym = (f & NB_YM) ? (cym + cc) / 2.0 : 0.0;
    const double wyp = (f & NB_YP) ? (cc + cyp) / 2.0 : 0.0;
    const double wzm = (f & NB_ZM) ? (czm + cc) / 2.0 : 0.0;
    const double wzp = (f & NB_ZP) ? (cc + czp) / 2.0 : 0.0;
    double ds = 0.0;
    if (f & NB_XM) ds += wxm;
    if (f & NB_XP) ds += wxp;
    if (f & NB_YM) ds += wym;
    if (f & NB_YP) ds += wyp;
    if (f & NB_ZM) ds += wzm;
    if (f & NB_ZP) ds += wzp;
    if (f & NB_ZM) acc = T::fma_r(wzm * g.sc, T::mulr(xzm, pre), acc);
    if (f & NB_YM) acc = T::fma_r(wym * g.sc, T::mulr(xym, pre), acc);
    if (f & NB_XM) acc = T::fma_r(wxm * g.sc, T::mulr(xxm, pre), acc);
    acc = T::fma_r(-ds * g.sc, xc, acc);
    if (f & NB_XP) acc = T::fma_r(wxp * g.sc, T::mulr(xxp, pre), acc);
    if (f & NB_YP) acc = T::fma_r(wyp * g.sc, T::mulr(xyp, pre), acc);
    if (f & NB_ZP) acc = T::fma_r(wzp * g.sc, T::mulr(xzp, pre), acc);
  }
  return acc;
}

// NB_* flags of a site from its flat index: the reference's triplet-loop conditions
// (laplacians.hpp:33-42, :133-148), evaluated once per site
__device__ __forceinline__ unsigned site_flags(const Geom &g, long long idx) {
  const long long gi = idx + g.goff;
  const unsigned ugi = (unsigned)gi;
  const unsigned i = ugi % (unsigned)g.nx;
  const unsigned jj = ugi / (unsigned)g.nx;
  unsigned f = 0;
  if (i != 0) f |= NB_XM;
  if (i != (unsigned)(g.nx - 1)) f |= NB_XP;
  if (gi >= g.nx) f |= NB_YM;
  if (gi + g.nx < g.nglob) f |= NB_YP;
  bool ring;
  if (g.dim3) {
    if (gi >= g.plane) f |= NB_ZM;
    if (gi + g.plane < g.nglob) f |= NB_ZP;
    const unsigned j = jj % (unsigned)g.ny, k = jj / (unsigned)g.ny;
    ring = i == 0 || i == (unsigned)(g.nx - 1) || j == 0 || j == (unsigned)(g.ny - 1) || k == 0 ||
           k == (unsigned)(g.nz_glob - 1);
  } else {
    ring = i == 0 || i == (unsigned)(g.nx - 1) || jj == 0 || jj == (unsigned)(g.ny - 1);
  }
  if (ring) f |= NB_RING;
  return f;
}

// Plane-marching work decomposition of a 3D slab: a CTA owns a kTileX x kTileY in-plane tile and
// marches it through kz planes; z+-1 neighbours come from registers, in-plane neighbours from L1
// (they were fetched as "next plane" values by the neighbouring threads one step earlier), so every
// stencil input is pulled from L2/HBM about once instead of five times.
constexpr int kTileX = 32, kTileY = 8; // kTileX * kTileY == kThreads
struct March {
  int ntx, nty, nzc, zc; // tiles in x, y; z-chunks; planes per chunk
  long long units;       // ntx * nty * nzc
};
__host__ __device__ inline March make_march(const Geom &g, int target_units) {
  March m;
  m.ntx = (g.nx + kTileX - 1) / kTileX;
  m.nty = (g.ny + kTileY - 1) / kTileY;
  const long long tiles = (long long)m.ntx * m.nty;
  long long chunks = (target_units + tiles - 1) / tiles;
  if (chunks < 1) chunks = 1;
  int zc = (int)((g.nz_loc + chunks - 1) / chunks);
  if (zc < 8) zc = 8;
  if (zc > g.nz_loc) zc = g.nz_loc;
  m.zc = zc;
  m.nzc = (g.nz_loc + zc - 1) / zc;
  m.units = tiles * m.nzc;
  return m;
}

template <class S> struct GlobalFetch {
  const S *__restrict__ p;
  __device__ __forceinline__ S operator()(long long i) const { return p[i]; }
};
template <class S, bool ANISO>
__device__ __forceinline__ S stencil_site(const Geom &g, const S *__restrict__ x,
                                          const double *__restrict__ cf, long long idx, double pre) {
  return stencil_site_f<S, ANISO>(g, GlobalFetch<S>{x}, cf, idx, pre);
}

// clamp a local flat index onto the interior: the fused form of neumann_bc_no_velocity
// (nlsolvers/host/include/boundaries.hpp:40-57, boundaries_3d.hpp:8-32).  The reference
// copies x-faces, then y-faces (all x), then z-faces (all x,y); the composition is
// u(i,j,k) <- u_new(clamp(i), clamp(j), clamp(k)) with clamp(a) = min(max(a,1), n_a-2).
__device__ __forceinline__ long long clamp_site(const Geom &g, long long idx) {
  const unsigned ugi = (unsigned)(idx + g.goff);
  unsigned i = ugi % (unsigned)g.nx;
  unsigned jj = ugi / (unsigned)g.nx;
  unsigned j = g.dim3 ? jj % (unsigned)g.ny : jj;
  unsigned k = g.dim3 ? jj / (unsigned)g.ny : 0u;
  i = min(max(i, 1u), (unsigned)(g.nx - 2));
  j = min(max(j, 1u), (unsigned)(g.ny - 2));
  if (g.dim3) k = min(max(k, 1u), (unsigned)(g.nz_glob - 2));
  const long long gi = ((long long)k * g.ny + j) * g.nx + i;
  return gi - g.goff; // may point into a ghost plane only if nz_loc == 1 (rejected at create)
}

// ---------------------------------------------------------------- reductions
// Deterministic two-stage grid reduction of K doubles per trajectory:
//   stage 1: warp shuffles -> shared memory -> one partial per CTA in `partials`
//   stage 2: the last CTA of the trajectory to arrive (atomic ticket) sums the partials in
//            a fixed order and stores the K totals to `dst`; no host round trip, no
//            floating-point atomics, bit-reproducible for a given launch shape.
// z-slab decomposition (one handle per GPU): cross-GPU sum of the per-slab totals through
// peer-mapped mailboxes, entirely inside the producing kernel (no host, no NCCL call):
//   the last CTA of a slab writes its K totals into slot [seq % kSlabSlots][rank] of EVERY
//   rank's mailbox (peer stores over NVLink), fences, then release-stores the sequence number
//   seq+1 into the matching flag; it then spins (acquire loads, LOCAL memory) until all G flags
//   of its own mailbox carry seq+1 and sums the G contributions in rank order, so every rank
//   obtains bit-identical scalars.  The same release/acquire chain orders the halo-plane peer
//   stores of that kernel before the next kernel's stencil reads on the neighbour.
constexpr int kSlabMaxRanks = 8;
constexpr int kSlabSlots = 4;
constexpr int kSlabMaxK = 32;
struct SlabMailbox { // lives at offset 0 of every rank's exchange arena
  double red[kSlabSlots][kSlabMaxRanks][kSlabMaxK];
  unsigned long long flag[kSlabSlots][kSlabMaxRanks];
  unsigned long long seq;   // reductions completed by the owning rank
  unsigned abort_flag;      // set once a wait timed out: later waits are skipped
};
struct SlabCtx {
  int rank, nranks;
  SlabMailbox *box[kSlabMaxRanks]; // box[rank] is local, the others are peer mappings
  int *status;
};

template <class S> struct Halo { // peer ghost planes of the vector a kernel produces (or null)
  S *lo; // lower neighbour's upper ghost plane
  S *hi; // upper neighbour's lower ghost plane
};
template <class S>
__device__ __forceinline__ void halo_push(const Halo<S> &h, long long n, long long plane, long long idx, S v) {
  if (h.lo != nullptr && idx < plane) h.lo[idx] = v;
  if (h.hi != nullptr && idx >= n - plane) h.hi[idx - (n - plane)] = v;
}

__device__ __forceinline__ void st_release_sys_u64(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

struct RedCtx {
  double *partials;  // [batch][gridDim.x][K]
  unsigned *tickets; // [batch]
  const SlabCtx *slab; // null unless z-slab decomposed (batch == 1 then)
};

// called by all threads of the last CTA; s_tot[K] holds this slab's totals
__device__ __forceinline__ void slab_allreduce_rt(const SlabCtx &sc, int K, double *s_tot, double *dst, double *dst1,
                                                  int split) {
  SlabMailbox *mine = sc.box[sc.rank];
  const unsigned long long seq = *(volatile unsigned long long *)&mine->seq; // written only by this rank's previous reduction
  const int slot = (int)(seq % kSlabSlots);
  const int G = sc.nranks;
  for (int t = threadIdx.x; t < G * K; t += blockDim.x) {
    const int q = t / K, k = t - q * K;
    sc.box[q]->red[slot][sc.rank][k] = s_tot[k];
  }
  __threadfence_system();
  __syncthreads();
  if ((int)threadIdx.x < G) st_release_sys_u64(&sc.box[threadIdx.x]->flag[slot][sc.rank], seq + 1);
  if ((int)threadIdx.x < G) {
    const long long t0 = clock64();
    bool dead = *(volatile unsigned *)&mine->abort_flag != 0u;
    while (!dead && ld_acquire_sys_u64(&mine->flag[slot][threadIdx.x]) != seq + 1) {
      if (clock64() - t0 > 6000000000LL) { // ~3 s: a peer died or never launched
        dead = true;
        mine->abort_flag = 1u;
        atomicExch(sc.status, 2);
      }
    }
  }
  __syncthreads();
  if ((int)threadIdx.x < K) {
    double v = 0.0;
    for (int q = 0; q < G; ++q) v += *(volatile double *)&mine->red[slot][q][threadIdx.x];
    if ((int)threadIdx.x < split) dst[threadIdx.x] = v;
    else dst1[threadIdx.x - split] = v;
  }
  if (threadIdx.x == 0) mine->seq = seq + 1;
}
template <int K>
__device__ __forceinline__ void slab_allreduce(const SlabCtx &sc, double *s_tot, double *dst, double *dst1, int split) {
  static_assert(K <= kSlabMaxK, "reduction too wide for the slab mailbox");
  slab_allreduce_rt(sc, K, s_tot, dst, dst1, split);
}

template <int K>
__device__ __forceinline__ void grid_reduce_impl(double (&acc)[K], const RedCtx &rc, double *dst, double *dst1, int split) {
  __shared__ double s_part[kThreads / 32][K];
  __shared__ double s_tot[K];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double v = acc[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) s_part[warp][k] = v;
  }
  __syncthreads();
  const int b = blockIdx.y, nblk = gridDim.x;
  double *mine = rc.partials + ((size_t)b * nblk + blockIdx.x) * K;
  if (threadIdx.x < K) {
    double v = 0.0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) v += s_part[w][threadIdx.x];
    mine[threadIdx.x] = v;
  }
  // slab mode: this CTA's halo-plane peer stores must be visible system-wide before the flag
  if (rc.slab != nullptr) __threadfence_system();
  else __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned t = atomicAdd(&rc.tickets[b], 1u);
    s_last = (t == (unsigned)nblk - 1u);
    if (s_last) rc.tickets[b] = 0u; // re-arm for the next kernel on this stream
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const double *all = rc.partials + (size_t)b * nblk * K;
  const bool slab = rc.slab != nullptr;
  for (int k = warp; k < K; k += kThreads / 32) {
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
    slab_allreduce<K>(*rc.slab, s_tot, dst, dst1, split);
  }
}

template <int K>
__device__ __forceinline__ void grid_reduce(double (&acc)[K], const RedCtx &rc, double *dst) {
  grid_reduce_impl<K>(acc, rc, dst, dst, K);
}
// H+1 values: the first H totals go to dst0[0..H), the last one to dst1[0]
template <int H>
__device__ __forceinline__ void grid_reduce2s(double (&acc)[H + 1], const RedCtx &rc, double *dst0, double *dst1) {
  grid_reduce_impl<H + 1>(acc, rc, dst0, dst1, H);
}

} // namespace nls
