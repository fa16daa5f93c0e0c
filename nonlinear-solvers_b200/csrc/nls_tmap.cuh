// nls_tmap.cuh -- pass A of a Lanczos iteration (stencil + recurrence + dots) with the stencil input staged through
// shared memory by TENSOR-MAP TMA loads (cp.async.bulk.tensor.3d, SASS UTMALDG), for 3D single-trajectory grids.
//
// The flat pass (k_lanczos_a) fetches the seven stencil inputs of a site through L1/L2: only the x neighbours hit
// L1, the y / z neighbours (and, for div(c grad), six more c values) come from L2 for EVERY site -- for a complex
// div(c grad) pass with 6 basis columns that is ~264 B/site through L2 for ~120 B/site of DRAM traffic, i.e. the pass
// sits at the L2 throughput cap (~12 TB/s) at its measured 5.0 TB/s of DRAM traffic, and the narrow early iterations
// (2 / 4 columns) at 2.4 / 3.3 TB/s (profiles/r02g_launches_c3_flat.csv).  Here a CTA owns a 64 x 8 in-plane tile and
// marches it through a chunk of z planes.  One elected thread issues, two planes ahead, a 3-D box load of
// (64+2) x (8+2) x 1 sites of the input column (and of c) into a ring of shared-memory plane slots; out-of-range halo
// elements are zero-filled by the TMA unit, so faces need no address logic.  Every input site is fetched from L2/HBM
// once per tile (+ its halo ring, x1.29), the neighbour values are shared-memory reads, and the 64-bit index
// arithmetic leaves the inner loop.  The basis columns for the dots and the store of w stream straight through
// registers as before; every thread handles TWO sites of a plane (rows ty and ty+4) so that twice as many column
// loads are in flight per barrier.  Measured history (C3 / C5, DESIGN.md section 3 has the table):
//   v1  one site per thread, 64 x 4 tiles: latency-bound (long-scoreboard 12 per issue), 2.2-2.4 TB/s -- slower than flat
//       (profiles/r02j_ncu_full_c3_pass_a_tm_v1_vs_flat.txt);
//   v2  two sites per thread, column loads before the mbarrier wait: 3.7-4.3 TB/s, issue-bound (215 M warp instructions);
//   v3  warp-uniform interior / boundary paths (interior warps carry no flag logic) + compile-time column roles (MODE):
//       119 M instructions, 4.5-5.5 TB/s with div(c grad), 5.8-6.6 TB/s uniform; C3 10.6 -> 9.4, C5 61.7 -> 47.7 ms/step
//       (profiles/r02l_ncu_full_c3_pass_a_tm_v3.txt, r02m_launches_c3.csv, r02p_launches_c5.csv): the default;
//   v4  knobs: L2 prefetch of the column values two planes ahead (default, -1.2 %), per-warp ring hand-back through
//       EMPTY mbarriers (PIPE; removes the CTA barrier stall but loses to its own shared-memory footprint: opt-in).
// The rows of the operator -- incl. the reference's quirks (ring diagonal, y-neighbour test on the flat index:
// laplacians.hpp:124-148, :181-210) -- are those of stencil_vals (nls_device.cuh), the flat kernel's arithmetic.
#pragma once
#include <cuda.h>

#include <type_traits>

#include "nls_tma.cuh"

namespace nls {

constexpr int kTmBX = 64, kTmBY = 8;  // tile; kThreads = 64 x 4 threads, two rows per thread
constexpr int kTmRows = 2, kTmTY = kTmBY / kTmRows;
// plane slots of the shared-memory ring.  4: planes k-1, k, k+1 in use + one in flight, the slot of plane k-1 is handed
// back by a __syncthreads per plane step.  6 ("pipe"): three planes in flight, every warp hands a plane back by arriving
// on the slot's EMPTY mbarrier and only the issuing thread waits for it, one step later -- no CTA-wide barrier per plane.
constexpr int kTmR = 4, kTmRPipe = 6, kTmRMax = 8;
constexpr int kTmPY = kTmBY + 2;
static_assert(kTmBX * kTmTY == kThreads, "tile / thread mapping");
// Box rows.  A box row must span a multiple of 32 BYTES on sm_100a: a 66-double (528 B) row is accepted by
// cuTensorMapEncodeTiled and then faults the TMA unit with "illegal instruction", a 132-double (1056 B) row works
// (tools/microbench/tmap_probe.cu, profiles/r02i_tmap_probe.log).  Complex rows are 66 sites = 132 doubles; real rows
// (real states, the c field) are widened to 68 doubles and start one site further left.
template <int W> struct TmRow {
  static constexpr int PX = W == 2 ? kTmBX + 2 : kTmBX + 4; // sites per box row
  static constexpr int XO = W == 2 ? 1 : 2;                 // box offset of the tile's first site
  static constexpr int kSites = PX * kTmPY;
  static constexpr int kPad = (kSites * W * 8 + 127) / 128 * 128 / (W * 8); // slot stride in sites: 128-byte multiple
};
constexpr int kTmHdrBytes = 256; // FULL mbarriers [8] + EMPTY mbarriers [8] + abort word in front of the slots

// dynamic shared memory of one CTA with `nslot` plane slots
inline size_t tm_smem_bytes(int W, bool aniso, int nslot) {
  const size_t xs = (size_t)nslot * (W == 2 ? TmRow<2>::kPad * 16 : TmRow<1>::kPad * 8);
  return kTmHdrBytes + xs + (aniso ? (size_t)nslot * TmRow<1>::kPad * 8 : 0);
}

struct TmGeom {
  int ntx, nty, nzc, zc; // tiles in x / y, z chunks, planes per chunk
  long long units;
  int zoff;              // tensor-map z coordinate of local plane 0 (1 when the arrays carry ghost planes)
  int pf;                // planes the basis-column values are prefetched ahead into L2 (0 = off)
};

__device__ __forceinline__ void mbar_inval(unsigned long long *bar) {
  asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void tma_load_3d(void *dst_smem, const CUtensorMap *tm, int c0, int c1, int c2, unsigned long long *bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
               : "memory");
}

// NB_* flags of site (i, jj, global plane kg): the reference's triplet-loop conditions (site_flags, nls_device.cuh)
__device__ __forceinline__ unsigned tm_flags(const Geom &g, int i, int jj, int kg) {
  unsigned f = 0;
  if (i != 0) f |= NB_XM;
  if (i != g.nx - 1) f |= NB_XP;
  if (!(kg == 0 && jj == 0)) f |= NB_YM;                    // gi >= nx        (laplacians.hpp:140)
  if (!(kg == g.nz_glob - 1 && jj == g.ny - 1)) f |= NB_YP; // gi + nx < nglob (no plane check: y-wrap quirk)
  if (kg > 0) f |= NB_ZM;
  if (kg < g.nz_glob - 1) f |= NB_ZP;
  if (i == 0 || i == g.nx - 1 || jj == 0 || jj == g.ny - 1 || kg == 0 || kg == g.nz_glob - 1) f |= NB_RING;
  return f;
}

// (L x)(site) * inv from the plane slots.  EDGE = false: every neighbour exists and none wraps (all flags set, no
// ring diagonal) -- the compiler folds the flag selects away; EDGE = true: flags of the site, and the wrap partners
// (two lines of sites per plane) fetched from global memory
template <class S, bool ANISO, bool EDGE>
__device__ __forceinline__ S tm_stencil(const Geom &g, unsigned f, double inv, const S *__restrict__ xm, const S *__restrict__ xc,
                                        const S *__restrict__ xp, const double *__restrict__ cm, const double *__restrict__ cc,
                                        const double *__restrict__ cp, int sidx, int cidx, int jj, long long idx,
                                        const S *__restrict__ cj, const double *__restrict__ cfb, S &centre) {
  constexpr int W = Sc<S>::W;
  using RX = TmRow<W>;
  using RC = TmRow<1>;
  centre = xc[sidx];
  S xym = xc[sidx - RX::PX], xyp = xc[sidx + RX::PX];
  const S xxm = xc[sidx - 1], xxp = xc[sidx + 1];
  const S xzm = xm[sidx], xzp = xp[sidx];
  double c0 = 0.0, czm = 0.0, czp = 0.0, cym = 0.0, cyp = 0.0, cxm = 0.0, cxp = 0.0;
  if (ANISO) {
    c0 = cc[cidx]; cym = cc[cidx - RC::PX]; cyp = cc[cidx + RC::PX];
    cxm = cc[cidx - 1]; cxp = cc[cidx + 1];
    czm = cm[cidx]; czp = cp[cidx];
  }
  if (!EDGE) {
    constexpr unsigned kAll = NB_XM | NB_XP | NB_YM | NB_YP | NB_ZM | NB_ZP;
    return stencil_vals<S, ANISO>(g, kAll, inv, centre, xzm, xym, xxm, xxp, xyp, xzp, c0, czm, cym, cxm, cxp, cyp, czp);
  }
  // the reference's y links ignore plane boundaries: (i, 0, k) <-> (i, ny-1, k-1).  Those partners are not in the box
  // (its out-of-range rows are zero-filled): fetch them directly
  if (jj == 0 && (f & NB_YM)) {
    xym = cj[idx - g.nx];
    if (ANISO) cym = cfb[idx - g.nx];
  }
  if (jj == g.ny - 1 && (f & NB_YP)) {
    xyp = cj[idx + g.nx];
    if (ANISO) cyp = cfb[idx + g.nx];
  }
  return stencil_vals<S, ANISO>(g, f, inv, centre, xzm, xym, xxm, xxp, xyp, xzp, c0, czm, cym, cxm, cxp, cyp, czp);
}

// Columns [i0, i0+NC) of 0..j take part in the dots; the kernel also forms w (first chunk of the iteration).
// The role of every chunk column is a compile-time fact, so the site loop has no selects between value sources:
//   TM_TAIL: the chunk ends with columns j-1 and j (i0 + NC - 1 == j, j >= 1): those two are at hand (the recurrence
//            term, the stencil centre), only the first NC-2 columns are loaded;
//   TM_BODY: the chunk holds neither (i0 + NC <= j - 1): all NC columns are loaded;
//   TM_SELF: the chunk is column j alone (NC == 1: j == 0, or the three-term recurrence without re-orthogonalisation).
enum { TM_TAIL = 0, TM_BODY = 1, TM_SELF = 2 };
template <class S, bool ANISO, int NC, int MODE, bool PIPE = false>
__global__ void __launch_bounds__(kThreads, 2)
k_lanczos_a_tm(const __grid_constant__ CUtensorMap tmx, const __grid_constant__ CUtensorMap tmc, Geom g, TmGeom tg,
               const S *__restrict__ start, S *__restrict__ V, size_t colstride, int j, int i0,
               const double *__restrict__ cf, double *scal, RedCtx rc, int *status) {
  using T = Sc<S>;
  constexpr int W = T::W;
  using RX = TmRow<W>;
  using RC = TmRow<1>;
  constexpr int kLd = MODE == TM_TAIL ? NC - 2 : (MODE == TM_BODY ? NC : 0); // columns loaded for the dots
  static_assert(MODE != TM_TAIL || NC >= 2, "tail chunk holds columns j-1 and j");
  static_assert(MODE != TM_SELF || NC == 1, "self chunk is column j alone");
  extern __shared__ __align__(128) unsigned char tm_smem[];
  unsigned long long *full = (unsigned long long *)tm_smem, *empty = full + kTmRMax;
  volatile int *s_abort = (volatile int *)(tm_smem + 128);
  constexpr int R = PIPE ? kTmRPipe : kTmR;
  S *xs = (S *)(tm_smem + kTmHdrBytes);                  // [R][RX::kPad]
  double *cs = (double *)(xs + (size_t)R * RX::kPad);    // [R][RC::kPad] (ANISO only)
  double *sb = scal;
  const S *cj = colptr<S>(start, V, colstride, j) + g.halo;
  const S *cm = j > 0 ? colptr<S>(start, V, colstride, j - 1) + g.halo : nullptr;
  S *wc = V + (size_t)j * colstride + g.halo; // column j+1
  const S *cb = colptr<S>(start, V, colstride, i0) + g.halo;
  const size_t cs0 = i0 == 0 ? (size_t)((V + 0) - start) : colstride;
  const double *cfb = ANISO ? cf + g.halo : nullptr;
  const double bj = sqrt(sb[kOffNrm2 + j]);
  const double inv = g.raw ? 1.0 : 1.0 / bj; // raw: w = L column_j as it is (NLS_REORTH_GRAM)
  const double cmul = (j > 0 && !g.raw) ? -bj / sqrt(sb[kOffNrm2 + j - 1]) : 0.0;
  double acc[NC * W];
#pragma unroll
  for (int k = 0; k < NC * W; ++k) acc[k] = 0.0;
  const int tx = threadIdx.x % kTmBX, ty = threadIdx.x / kTmBX;
  constexpr unsigned kBoxBytes = (unsigned)(RX::kSites * sizeof(S)) + (ANISO ? (unsigned)(RC::kSites * 8) : 0u);
  if (threadIdx.x == 0) *s_abort = 0;
  const long long tiles = (long long)tg.ntx * tg.nty;
  bool first_unit = true;
  for (long long unit = blockIdx.x; unit < tg.units; unit += gridDim.x) {
    const int tile = (int)(unit % tiles), zc = (int)(unit / tiles);
    const int x0 = (tile % tg.ntx) * kTmBX, y0 = (tile / tg.ntx) * kTmBY;
    const int k0 = zc * tg.zc, k1 = min(k0 + tg.zc, g.nz_loc);
    const int np = k1 - k0 + 2; // planes of this unit: ring plane p <-> local plane k0 - 1 + p, slot p % R, use p / R
    const int i = x0 + tx;
    const int jj[kTmRows] = {y0 + ty, y0 + ty + kTmTY};
    const bool live[kTmRows] = {i < g.nx && jj[0] < g.ny, i < g.nx && jj[1] < g.ny};
    // warp-uniform: some site of this warp lies on an x / y face of the grid (or outside it)
    const bool edge_xy = __any_sync(0xffffffffu, !live[1] || i == 0 || i == g.nx - 1 || jj[0] == 0 || jj[1] == g.ny - 1 ||
                                                     jj[0] == g.ny - 1 || jj[1] == 0);
    // box origin in tensor-map coordinates: the inner dimension counts doubles (complex = 2 per site)
    const int bx = (x0 - RX::XO) * W, by = y0 - 1;
    auto issue = [&](int p, int slot) {
      mbar_expect_tx(&full[slot], kBoxBytes);
      tma_load_3d(xs + (size_t)slot * RX::kPad, &tmx, bx, by, k0 - 1 + p + tg.zoff, &full[slot]);
      if (ANISO) tma_load_3d(cs + (size_t)slot * RC::kPad, &tmc, x0 - RC::XO, by, k0 - 1 + p + tg.zoff, &full[slot]);
    };
    __syncthreads(); // the previous unit's reads of the ring and its arrivals are done
    if (threadIdx.x == 0) { // fresh barriers per unit: the parity of plane p's completion is (p / R) & 1, no bookkeeping
      for (int r = 0; r < R; ++r) {
        if (!first_unit) { mbar_inval(&full[r]); mbar_inval(&empty[r]); }
        mbar_init(&full[r], 1);
        mbar_init(&empty[r], kThreads / 32);
      }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    first_unit = false;
    __syncthreads();
    constexpr int lead = PIPE ? R - 1 : 4; // planes issued up front
    if (threadIdx.x == 0) {
      fence_proxy_async();
      for (int p = 0; p < min(lead, np); ++p) issue(p, p);
    }
    mbar_wait(&full[0], 0u, s_abort, status); // planes k0-1 and k0
    mbar_wait(&full[1], 0u, s_abort, status);
    const int sidx[kTmRows] = {(ty + 1) * RX::PX + tx + RX::XO, (ty + kTmTY + 1) * RX::PX + tx + RX::XO};
    const int cidx[kTmRows] = {(ty + 1) * RC::PX + tx + RC::XO, (ty + kTmTY + 1) * RC::PX + tx + RC::XO};
    long long idx0 = ((long long)k0 * g.ny + jj[0]) * g.nx + i;
    const long long rowskip = (long long)kTmTY * g.nx;
    for (int k = k0; k < k1; ++k, idx0 += g.plane) {
      const int pc = k - k0 + 1; // ring plane of the centre; slots of planes pc-1, pc, pc+1:
      const int sm = (pc - 1) % R, sc = pc % R, sp = (pc + 1) % R;
      const unsigned par_p = (unsigned)((pc + 1) / R) & 1u; // parity of plane pc+1's completion
      const int kg = g.z0 + k;
      const bool edge = edge_xy || kg == 0 || kg == g.nz_glob - 1; // warp-uniform
      const S *xsm = xs + (size_t)sm * RX::kPad, *xsc = xs + (size_t)sc * RX::kPad, *xsp = xs + (size_t)sp * RX::kPad;
      const double *csm = cs + (size_t)sm * RC::kPad, *csc = cs + (size_t)sc * RC::kPad, *csp = cs + (size_t)sp * RC::kPad;
      // one plane step of both sites: the global loads first (independent of the ring), then the wait for plane k+1
      auto step = [&](auto edge_c) {
        constexpr bool EDGE = decltype(edge_c)::value;
        S cmv[kTmRows], v[kTmRows][kLd > 0 ? kLd : 1];
#pragma unroll
        for (int r = 0; r < kTmRows; ++r) {
          const long long idx = idx0 + r * rowskip;
          const bool on = !EDGE || live[r];
          cmv[r] = T::zero();
          if (on && j > 0) cmv[r] = ld_stream(cm + idx);
#pragma unroll
          for (int q = 0; q < kLd; ++q)
            if (on) v[r][q] = ld_stream((q == 0 ? cb : cb + cs0 + (size_t)(q - 1) * colstride) + idx);
          if (tg.pf && on && k + tg.pf < k1) { // the same values of a later plane -> L2
            const long long ipf = idx + (long long)tg.pf * g.plane;
            if (j > 0) pf_l2(cm + ipf);
#pragma unroll
            for (int q = 0; q < kLd; ++q) pf_l2((q == 0 ? cb : cb + cs0 + (size_t)(q - 1) * colstride) + ipf);
          }
        }
        mbar_wait(&full[sp], par_p, s_abort, status);
#pragma unroll
        for (int r = 0; r < kTmRows; ++r) {
          const long long idx = idx0 + r * rowskip;
          if (EDGE && !live[r]) continue;
          const unsigned f = EDGE ? tm_flags(g, i, jj[r], kg) : 0u;
          S xc;
          S w = tm_stencil<S, ANISO, EDGE>(g, f, inv, xsm, xsc, xsp, csm, csc, csp, sidx[r], cidx[r], jj[r], idx, cj, cfb, xc);
          if (j > 0) w = T::fma_r(cmul, cmv[r], w);
          __stcs(wc + idx, w);
#pragma unroll
          for (int q = 0; q < kLd; ++q) T::dot_acc(acc + q * W, v[r][q], w);
          if constexpr (MODE == TM_TAIL) T::dot_acc(acc + (NC - 2) * W, cmv[r], w);
          if constexpr (MODE != TM_BODY) T::dot_acc(acc + (NC - 1) * W, xc, w);
        }
      };
      if (edge) step(std::true_type{});
      else step(std::false_type{});
      if constexpr (PIPE) {
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(&empty[sm]); // this warp is done with plane pc-1
        if (threadIdx.x == 0 && pc + R - 2 < np) {            // plane pc+R-2 goes where plane pc-2 was
          if (pc >= 2) mbar_wait(&empty[(pc - 2) % R], (unsigned)((pc - 2) / R) & 1u, s_abort, status); // handed back one step ago
          fence_proxy_async();
          issue(pc + R - 2, (pc + R - 2) % R);
        }
      } else {
        __syncthreads(); // all reads of plane pc-1 are done: plane pc+3 goes there, two steps ahead of its use
        if (threadIdx.x == 0 && pc + 3 < np) {
          fence_proxy_async();
          issue(pc + 3, (pc + 3) % R);
        }
      }
    }
  }
  grid_reduce<NC * W>(acc, rc, sb + kOffH + i0 * W);
}

} // namespace nls
