// nls_tmap.cuh -- pass A of a Lanczos iteration (stencil + recurrence + dots) with the stencil input staged through
// shared memory by TENSOR-MAP TMA loads (cp.async.bulk.tensor.3d, SASS UTMALDG), for 3D single-trajectory grids.
//
// The flat pass (k_lanczos_a) fetches the seven stencil inputs of a site through L1/L2: only the x neighbours hit
// L1, the y / z neighbours (and, for div(c grad), six more c values) come from L2 for EVERY site -- 5 x 16 B (+ 5 x 8 B)
// of L2 traffic per site on top of the streamed basis columns, which is what holds the stencil-carrying launches at
// 2.4-4.3 TB/s (profiles/r01d_launches_c3_dram.csv).  Here a CTA owns a BX x BY in-plane tile and marches it through a
// chunk of z planes.  One elected thread issues, two planes ahead, a 3-D box load of (BX+2) x (BY+2) x 1 sites of the
// input column (and of c) into a ring of shared-memory plane slots; out-of-range halo elements are zero-filled by the
// TMA unit, so faces need no address logic.  Every input site is fetched from L2/HBM once per tile (+ its halo ring),
// the neighbour values are shared-memory reads, and the 64-bit index arithmetic leaves the inner loop.  The basis
// columns for the dots and the store of w stream straight through registers as before.  The rows of the operator --
// incl. the reference's quirks (ring diagonal, y-neighbour test on the flat index: laplacians.hpp:124-148, :181-210) --
// are those of stencil_vals (nls_device.cuh), bit for bit the flat kernel's arithmetic.
#pragma once
#include <cuda.h>

#include "nls_tma.cuh"

namespace nls {

constexpr int kTmBX = 64, kTmBY = 4;  // tile: one site per thread and plane (kTmBX * kTmBY == kThreads)
constexpr int kTmR = 4;               // plane slots: k-1, k, k+1 in use + one in flight (plane k+3 re-uses k-1's slot)
constexpr int kTmPY = kTmBY + 2;
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

struct TmGeom {
  int ntx, nty, nzc, zc; // tiles in x / y, z chunks, planes per chunk
  long long units;
  int zoff;              // tensor-map z coordinate of local plane 0 (1 when the arrays carry ghost planes)
};

__device__ __forceinline__ void tma_load_3d(void *dst_smem, const CUtensorMap *tm, int c0, int c1, int c2, unsigned long long *bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
               : "memory");
}

// columns [i0, i0+cnt) of 0..j take part in the dots (first chunk of the iteration: also forms w)
template <class S, bool ANISO, int NC>
__global__ void __launch_bounds__(kThreads, 2)
k_lanczos_a_tm(const __grid_constant__ CUtensorMap tmx, const __grid_constant__ CUtensorMap tmc, Geom g, TmGeom tg,
               const S *__restrict__ start, S *__restrict__ V, size_t colstride, int j, int i0, int cnt,
               const double *__restrict__ cf, double *scal, RedCtx rc, int *status) {
  using T = Sc<S>;
  constexpr int W = T::W;
  using RX = TmRow<W>;
  using RC = TmRow<1>;
  __shared__ __align__(128) S xs[kTmR][RX::kPad];
  __shared__ __align__(128) double cs[ANISO ? kTmR : 1][ANISO ? RC::kPad : 16];
  __shared__ __align__(8) unsigned long long bars[kTmR];
  __shared__ int s_abort;
  double *sb = scal;
  const S *cj = colptr<S>(start, V, colstride, j) + g.halo;
  const S *cm = j > 0 ? colptr<S>(start, V, colstride, j - 1) + g.halo : nullptr;
  S *wc = V + (size_t)j * colstride + g.halo; // column j+1
  const S *cb = colptr<S>(start, V, colstride, i0) + g.halo;
  const size_t cs0 = i0 == 0 ? (size_t)((V + 0) - start) : colstride;
  const double *cfb = ANISO ? cf + g.halo : nullptr;
  const double bj = sqrt(sb[kOffNrm2 + j]);
  const double inv = 1.0 / bj;
  const double cmul = j > 0 ? -bj / sqrt(sb[kOffNrm2 + j - 1]) : 0.0;
  double acc[NC * W];
#pragma unroll
  for (int k = 0; k < NC * W; ++k) acc[k] = 0.0;
  const int tx = threadIdx.x % kTmBX, ty = threadIdx.x / kTmBX;
  constexpr unsigned kBoxBytes = (unsigned)(RX::kSites * sizeof(S)) + (ANISO ? (unsigned)(RC::kSites * 8) : 0u);
  if (threadIdx.x == 0) {
    for (int r = 0; r < kTmR; ++r) mbar_init(&bars[r], 1);
    s_abort = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  unsigned phase_bits = 0; // parity of every slot's next completion
  const long long tiles = (long long)tg.ntx * tg.nty;
  for (long long unit = blockIdx.x; unit < tg.units; unit += gridDim.x) {
    const int tile = (int)(unit % tiles), zc = (int)(unit / tiles);
    const int x0 = (tile % tg.ntx) * kTmBX, y0 = (tile / tg.ntx) * kTmBY;
    const int k0 = zc * tg.zc, k1 = min(k0 + tg.zc, g.nz_loc);
    const int i = x0 + tx, jj = y0 + ty;
    const bool live = i < g.nx && jj < g.ny;
    // box origin in tensor-map coordinates: the inner dimension counts doubles (complex = 2 per site)
    const int bx = (x0 - RX::XO) * W, by = y0 - 1;
    auto issue = [&](int k) { // plane k (local index, may be k0-1 .. k1) into slot (k - k0 + 1) % R
      const int slot = (k - k0 + 1) % kTmR;
      mbar_expect_tx(&bars[slot], kBoxBytes);
      tma_load_3d(&xs[slot][0], &tmx, bx, by, k + tg.zoff, &bars[slot]);
      if (ANISO) tma_load_3d(&cs[slot][0], &tmc, x0 - RC::XO, by, k + tg.zoff, &bars[slot]);
    };
    __syncthreads(); // the previous unit's reads of the ring are done
    if (threadIdx.x == 0) {
      fence_proxy_async();
      for (int k = k0 - 1; k <= min(k0 + 2, k1); ++k) issue(k);
    }
    // wait for planes k0-1 and k0 (slots 0, 1)
    for (int q = 0; q < 2; ++q) {
      mbar_wait(&bars[q], (phase_bits >> q) & 1u, &s_abort, status);
      phase_bits ^= 1u << q;
    }
    unsigned fxy = 0;
    if (i != 0) fxy |= NB_XM;
    if (i != g.nx - 1) fxy |= NB_XP;
    const bool ring_xy = i == 0 || i == g.nx - 1 || jj == 0 || jj == g.ny - 1;
    const int sidx = (ty + 1) * RX::PX + tx + RX::XO; // this thread's site inside a slot
    const int cidx = (ty + 1) * RC::PX + tx + RC::XO; // ... and inside a c slot
    long long idx = ((long long)k0 * g.ny + jj) * g.nx + i;
    for (int k = k0; k < k1; ++k, idx += g.plane) {
      const int sm = (k - k0) % kTmR, sc = (k - k0 + 1) % kTmR, sp = (k - k0 + 2) % kTmR;
      // plane k+1 must have landed
      mbar_wait(&bars[sp], (phase_bits >> sp) & 1u, &s_abort, status);
      phase_bits ^= 1u << sp;
      if (live) {
        const int kg = g.z0 + k;
        unsigned f = fxy;
        if (!(kg == 0 && jj == 0)) f |= NB_YM;                    // gi >= nx        (laplacians.hpp:140)
        if (!(kg == g.nz_glob - 1 && jj == g.ny - 1)) f |= NB_YP; // gi + nx < nglob (no plane check: y-wrap quirk)
        if (kg > 0) f |= NB_ZM;
        if (kg < g.nz_glob - 1) f |= NB_ZP;
        if (ring_xy || kg == 0 || kg == g.nz_glob - 1) f |= NB_RING;
        const S xc = xs[sc][sidx];
        S xym = xs[sc][sidx - RX::PX], xyp = xs[sc][sidx + RX::PX];
        const S xxm = xs[sc][sidx - 1], xxp = xs[sc][sidx + 1];
        const S xzm = xs[sm][sidx], xzp = xs[sp][sidx];
        double cc = 0.0, czm = 0.0, czp = 0.0, cym = 0.0, cyp = 0.0, cxm = 0.0, cxp = 0.0;
        if (ANISO) {
          cc = cs[sc][cidx]; cym = cs[sc][cidx - RC::PX]; cyp = cs[sc][cidx + RC::PX];
          cxm = cs[sc][cidx - 1]; cxp = cs[sc][cidx + 1];
          czm = cs[sm][cidx]; czp = cs[sp][cidx];
        }
        // the reference's y links ignore plane boundaries: (i, 0, k) <-> (i, ny-1, k-1).  Those partners are not in
        // the box (its out-of-range rows are zero-filled): fetch them directly, two lines of sites per plane
        if (jj == 0 && (f & NB_YM)) {
          xym = cj[idx - g.nx];
          if (ANISO) cym = cfb[idx - g.nx];
        }
        if (jj == g.ny - 1 && (f & NB_YP)) {
          xyp = cj[idx + g.nx];
          if (ANISO) cyp = cfb[idx + g.nx];
        }
        S w = stencil_vals<S, ANISO>(g, f, inv, xc, xzm, xym, xxm, xxp, xyp, xzp, cc, czm, cym, cxm, cxp, cyp, czp);
        S cmv = T::zero();
        if (j > 0) {
          cmv = ld_stream(cm + idx);
          w = T::fma_r(cmul, cmv, w);
        }
        __stcs(wc + idx, w);
#pragma unroll
        for (int q = 0; q < NC; ++q)
          if (q < cnt) {
            const int col = i0 + q;
            S vq;
            if (col == j) vq = xc;
            else if (col == j - 1) vq = cmv;
            else vq = ld_stream((q == 0 ? cb : cb + cs0 + (size_t)(q - 1) * colstride) + idx);
            T::dot_acc(acc + q * W, vq, w);
          }
      }
      __syncthreads(); // all reads of slot sm (plane k-1) are done: plane k+3 goes there, two steps ahead of its use
      if (threadIdx.x == 0 && k + 3 <= k1) {
        fence_proxy_async();
        issue(k + 3);
      }
    }
  }
  grid_reduce<NC * W>(acc, rc, sb + kOffH + i0 * W);
}

} // namespace nls
