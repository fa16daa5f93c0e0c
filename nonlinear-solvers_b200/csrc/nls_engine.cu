// nls_engine.cu -- host side of libnls_b200.so: the C ABI of include/nls.h.
// Owns device memory, one stream per handle, CUDA-graph capture of a whole time step.
// No CPU fallback anywhere: without a CUDA device every entry point returns NLS_ERR_CUDA.
#include "nls_persist.cuh"
#include "nls_tma.cuh"
#include "nls_tmap.cuh"

#include <unistd.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

using namespace nls;

namespace {
constexpr long long kPkMinSites = 64; // persistent kernel: minimum sites per CTA (pk_setup)
// tensor-map pass A: per-warp ring hand-back (measured slower: C3 9.37 -> 10.03 ms/step, profiles/r02n_*; opt-in) and the
// L2 prefetch lead of the column values in planes (2: C3 9.37 -> 9.26, C5 47.46 -> 46.84)
constexpr int kTmDefaultPipe = 0, kTmDefaultPrefetch = 2;
constexpr int kDefaultPrefetch = 1; // L2 prefetch lead of the flat Lanczos passes, in grid sweeps (Geom::pf)
constexpr size_t kDefaultColPad = 0; // see col_pad_bytes()
thread_local std::string g_err;

int fail(int code, const std::string &msg) {
  g_err = msg;
  return code;
}

#define CU(call)                                                                          \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess)                                                                \
      return fail(NLS_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));      \
  } while (0)

template <class S> struct Ptrs {};
} // namespace

struct nls_engine {
  nls_desc d{};
  Geom g{};
  bool cplx = false;
  int W = 1;
  size_t esz = 8;
  int m = 10;
  int batch = 1;
  int nblk = 1; // capacity of the partial-sum array (max CTAs per trajectory)
  int nsm = 148;
  int smem_optin = 0;  // max dynamic shared memory per CTA
  bool gdr_auto = true; // NLS_REORTH_FULL runs the Gram-delayed form on the multi-kernel path (nls_create)
  int first_cap = 0;   // complex pass A: columns in the stencil-carrying first chunk (0 = rule in lanczos_t)
  bool tma_ok = false; // TMA-staged Lanczos passes usable (alignment + size), see tma_setup
  bool tm_ok = false;  // tensor-map pass A (nls_tmap.cuh) usable
  TmGeom tmg{};
  std::unordered_map<const void *, CUtensorMap> tmaps; // one 3-D tensor map per stencil-input vector (by base address)
  uint64_t tm_launches = 0;
  cudaStream_t stream = nullptr;
  // device memory (all n-vectors: [batch][stride] scalars)
  void *V = nullptr;     // m-1 columns
  void *wbuf[2] = {nullptr, nullptr};
  void *rs = nullptr;   // dense output of nls_get_field_resampled (allocated on first use)
  double2 *dft_wx = nullptr, *dft_wy = nullptr, *dft_t1 = nullptr; // cropped-DFT weights / intermediate (nls_get_field_fft_resampled)
  uint32_t dft_tx = 0, dft_ty = 0;
  size_t rs_bytes = 0;
  void *w0 = nullptr;    // start vector (rho / g / matfunc input)
  void *u[3] = {nullptr, nullptr, nullptr};
  void *acc = nullptr;   // Gautschi accumulator
  double *mf = nullptr, *cf = nullptr;
  double *scalA = nullptr, *scalB = nullptr, *coef = nullptr, *redout = nullptr;
  RedCtx rc{nullptr, nullptr, nullptr};
  int *status = nullptr;
  int poison = 0;        // first non-zero device status seen: the handle's state is unusable from then on (sticky)
  size_t colstride = 0;  // scalars between basis columns
  // stepping state
  int nphase = 1, phase = 0;
  bool start_valid = false;
  bool sewi_started = false; // NLS_INT_SEWI: u- is valid (the SS2 start-up step has been taken)
  double start_dt = 0.0;
  cudaGraphExec_t gexec[3] = {nullptr, nullptr, nullptr};
  double gdt[3] = {0, 0, 0};
  uint64_t glaunch[3] = {0, 0, 0}, gtma[3] = {0, 0, 0};
  uint64_t launches = 0, tma_launches = 0, cap_tma = 0;
  bool capturing = false;
  uint64_t cap_count = 0;
  // persistent cooperative kernel (nls_persist.cuh)
  bool pk_ok = false;
  int pk_ncta = 0;
  long long pk_csz = 0;
  size_t pk_smem = 0;
  double *pk_part = nullptr;
  unsigned *pk_bar = nullptr;
  uint64_t pk_launches = 0;
  // asynchronous transfers (nls_get_field_async / nls_set_field_async): a 2-deep device-side snapshot ring drained
  // by a D2H stream, a 2-deep staging ring filled by an H2D stream; the compute stream only does D2D copies
  cudaStream_t d2h_stream = nullptr, h2d_stream = nullptr;
  void *ring[2] = {nullptr, nullptr}, *stage[2] = {nullptr, nullptr};
  size_t ring_bytes[2] = {0, 0}, stage_bytes[2] = {0, 0};
  cudaEvent_t ring_ready[2] = {nullptr, nullptr}, ring_done[2] = {nullptr, nullptr};
  cudaEvent_t stage_ready[2] = {nullptr, nullptr}, stage_free[2] = {nullptr, nullptr};
  int ring_next = 0, stage_next = 0;
  bool async_used = false;
  // temporally blocked 2D Stoermer-Verlet kernel (k_pk_sv2)
  bool sv2_ok = false;
  Sv2Plan sv2{};
  unsigned long long *sv2_prog = nullptr; // [ncta] progress counters, then one abort word
  // z-slabs: the stencil-input vectors (+ the mailbox) live in one exchange arena that the
  // other ranks map through CUDA IPC (or address directly when they live in this process)
  bool slab = false;
  char *arena = nullptr;
  size_t slot_bytes = 0;      // bytes per arena slot on THIS rank
  int rank = 0, nranks = 1;
  bool connected = false;
  SlabCtx *slab_dev = nullptr;
  double *slab_scratch = nullptr;
  struct Peer { char *base = nullptr; bool ipc = false; uint64_t stride = 0; uint32_t nz = 0; };
  Peer peers[kSlabMaxRanks];

  size_t vec_elems() const { return (size_t)batch * g.stride; }
  size_t vec_bytes() const { return vec_elems() * esz; }
  dim3 grid() const { return dim3((unsigned)nblk, (unsigned)batch); }
  void count(bool tma = false) {
    if (capturing) { cap_count++; cap_tma += tma; }
    else { launches++; tma_launches += tma; }
  }
  void *sv(int p) const { return p == 0 ? w0 : wbuf[0]; } // NLSE start vector (rho) of u[p]
};

namespace {

// Persistent-style grids: (resident CTAs per SM for THIS kernel) x (number of SMs) CTAs per
// launch, split over the batch, so every launch is exactly one wave (a fixed 4 CTAs/SM grid ran
// 1.33 waves for the kernels that need > 64 registers).  Grid-stride loops cover the rest.
dim3 grid_for(nls_engine *e, const void *fn) {
  static thread_local std::unordered_map<const void *, int> occ_cache;
  auto it = occ_cache.find(fn);
  int occ;
  if (it == occ_cache.end()) {
    occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, kThreads, 0) != cudaSuccess || occ < 1) {
      cudaGetLastError();
      occ = 1;
    }
    occ_cache[fn] = occ;
  } else occ = it->second;
  long long per = ((long long)e->nsm * occ) / e->batch;
  if (per < 1) per = 1;
  const long long want = (e->g.n + kThreads - 1) / kThreads;
  if (per > want) per = want;
  if (per > e->nblk) per = e->nblk;
  return dim3((unsigned)per, (unsigned)e->batch);
}

#define LAUNCH(e, kern, gr, ...)                                                   \
  do {                                                                             \
    kern<<<grid_for(e, (const void *)kern), kThreads, 0, (e)->stream>>>(__VA_ARGS__); \
    (e)->count();                                                                  \
  } while (0)

template <class S>
S *col(nls_engine *e, int j) {
  return (S *)e->V + (size_t)j * e->colstride;
}

// Bytes added between consecutive basis columns (and arena slots).  Without it every column of a 256^3 or
// 1024 x 1024 x 128 complex grid starts exactly 2^28 / 2^31 bytes after the previous one, so the up to 13 streams a
// Lanczos pass reads concurrently alias onto the same DRAM banks (launch list profiles/r02g_launches_c3_flat.csv:
// pass B 6.6 TB/s with 8 columns, 4.0 TB/s with 12).  NLS_COL_PAD (bytes, multiple of 256) overrides for tuning.
size_t col_pad_bytes() {
  static const size_t pad = []() -> size_t {
    if (const char *s = getenv("NLS_COL_PAD")) return (size_t)strtoull(s, nullptr, 0) / 256 * 256;
    return kDefaultColPad;
  }();
  return pad;
}
size_t slot_bytes_of(size_t stride_elems, size_t esz) { return (stride_elems * esz + 255) / 256 * 256 + col_pad_bytes(); }

// ---------------------------------------------------------------- z-slab exchange arena
constexpr int kArenaSlots = 7; // wbuf0, wbuf1, w0, u0, u1, u2, cf; then the m-1 basis columns
constexpr size_t kMailboxBytes = (sizeof(SlabMailbox) + 4095) / 4096 * 4096;

struct SlabDesc { // the 128-byte blob of nls_slab_export
  cudaIpcMemHandle_t ipc; // 64 bytes
  uint64_t magic, pid, raw, stride;
  uint32_t nz, z0, nx, ny;
  int32_t device, esz;
  uint64_t pad[1];
};
static_assert(sizeof(SlabDesc) == 128, "slab descriptor must be 128 bytes");
constexpr uint64_t kSlabMagic = 0x4e4c53534c414231ull; // "NLSSLAB1"

// peer ghost planes of a local arena vector (null/null when not decomposed or not in the arena)
template <class S>
Halo<S> halo_for(nls_engine *e, const void *local) {
  Halo<S> h{nullptr, nullptr};
  if (!e->slab || !e->connected || !local) return h;
  const ptrdiff_t off = (const char *)local - (e->arena + kMailboxBytes);
  if (off < 0 || (size_t)off >= (size_t)(kArenaSlots + e->m - 1) * e->slot_bytes || off % (ptrdiff_t)e->slot_bytes)
    return h;
  const int slot = (int)(off / (ptrdiff_t)e->slot_bytes);
  const size_t plane = (size_t)e->g.plane;
  if (e->rank > 0) {
    const nls_engine::Peer &p = e->peers[e->rank - 1];
    const size_t pslot = slot_bytes_of(p.stride, e->esz);
    h.lo = (S *)(p.base + kMailboxBytes + slot * pslot) + (p.stride - plane); // its upper ghost plane
  }
  if (e->rank + 1 < e->nranks) {
    const nls_engine::Peer &p = e->peers[e->rank + 1];
    const size_t pslot = slot_bytes_of(p.stride, e->esz);
    h.hi = (S *)(p.base + kMailboxBytes + slot * pslot); // its lower ghost plane
  }
  return h;
}

// push the boundary planes of an uploaded stencil input to the neighbours (collective)
template <class S>
void halo_exchange(nls_engine *e, const void *buf) {
  if (!e->slab || !e->connected) return;
  k_halo_exchange<S><<<e->grid(), kThreads, 0, e->stream>>>(e->g, (const S *)buf, e->rc, halo_for<S>(e, buf),
                                                         e->slab_scratch);
  e->count();
}

// ---------------------------------------------------------------- Lanczos (K2)
// ---------------------------------------------------------------- TMA-staged passes (nls_tma.cuh)
bool tma_attr(const void *fn, int bytes) { // opt in to large dynamic shared memory once per (device, kernel)
  static thread_local std::unordered_map<uint64_t, bool> done_map;
  int dev = 0;
  cudaGetDevice(&dev);
  const uint64_t key = (uint64_t)(uintptr_t)fn * 64u + (uint64_t)dev; // function addresses are >= 64-byte apart
  auto &done = done_map;
  auto it = done.find(key);
  if (it != done.end()) return it->second;
  const bool ok = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) == cudaSuccess;
  if (!ok) cudaGetLastError();
  done[key] = ok;
  return ok;
}


// Bulk copies need 16-byte aligned addresses and sizes: always true for complex128 vectors; float64
// vectors (and the c field) need even nx, plane, n and strides.
void tma_setup(nls_engine *e) {
  const Geom &g = e->g;
  e->tma_ok = false;
  // opt-in: on B200 the register-staged passes are faster (256^3 complex, 12 columns: pass B 4.1 vs 3.8-6.6 TB/s,
  // pass A 0.6-2.2 vs 2.1-4.5 TB/s; profiles/r01e_launches_c3_tma.csv) -- the per-tile overhead of the
  // 256-site tiles that fit a multi-stage ring beside 12 columns is not yet hidden
  if (!(e->d.flags & NLS_FLAG_TMA) || e->batch != 1 || e->smem_optin < 160 * 1024) return;
  const bool even = g.nx % 2 == 0 && g.n % 2 == 0 && g.halo % 2 == 0 && g.stride % 2 == 0 &&
                    (!g.dim3 || g.plane % 2 == 0) && e->colstride % 2 == 0;
  if (!e->cplx && !even) return;
  if (e->cplx && e->d.laplacian == NLS_LAP_ANISO && !even) return;
  // opt in to the large dynamic shared memory now: not allowed later, inside a stream capture
  bool ok;
  if (e->cplx)
    ok = tma_attr((const void *)k_lanczos_a_tma<double2, false>, e->smem_optin) &&
         tma_attr((const void *)k_lanczos_a_tma<double2, true>, e->smem_optin) &&
         tma_attr((const void *)k_lanczos_b_tma<double2, false>, e->smem_optin) &&
         tma_attr((const void *)k_lanczos_b_tma<double2, true>, e->smem_optin);
  else
    ok = tma_attr((const void *)k_lanczos_a_tma<double, false>, e->smem_optin) &&
         tma_attr((const void *)k_lanczos_a_tma<double, true>, e->smem_optin) &&
         tma_attr((const void *)k_lanczos_b_tma<double, false>, e->smem_optin) &&
         tma_attr((const void *)k_lanczos_b_tma<double, true>, e->smem_optin);
  e->tma_ok = ok;
}

TmaPlan tma_plan(nls_engine *e, int nseg, size_t extra) {
  TmaPlan pl;
  pl.tile = kTmaTileBytes / (int)e->esz;
  pl.ntile = (int)((e->g.n + pl.tile - 1) / pl.tile);
  pl.stage_bytes = ((size_t)nseg * kTmaTileBytes + extra + 127) / 128 * 128;
  long long ns = ((long long)e->smem_optin - kTmaHdrBytes) / (long long)pl.stage_bytes;
  pl.nstages = (int)(ns > 4 ? 4 : ns);
  pl.smem_bytes = kTmaHdrBytes + (size_t)pl.nstages * pl.stage_bytes;
  return pl;
}

template <class S, bool ANISO>
bool launch_a_tma(nls_engine *e, const S *start, S *V, int j, double *scal) {
  if (!e->tma_ok || j + 1 > kTmaMaxCols || (e->d.reorth != NLS_REORTH_FULL && e->d.reorth != NLS_REORTH_BLOCK)) return false;
  const TmaPlan pl = tma_plan(e, j + 1 + (ANISO ? 10 : 5), 64);
  const void *fn = (const void *)k_lanczos_a_tma<S, ANISO>;
  if (pl.nstages < 2 || !tma_attr(fn, e->smem_optin)) return false;
  const int gx = pl.ntile < e->nsm ? pl.ntile : e->nsm;
  k_lanczos_a_tma<S, ANISO><<<dim3(gx, 1), kTmaThreads, pl.smem_bytes, e->stream>>>(e->g, pl, start, V, e->colstride, j,
                                                                                 e->cf, scal, e->rc, e->status);
  e->count(true);
  return true;
}

template <class S>
bool launch_b_tma(nls_engine *e, const S *start, S *V, int j, int i0, int cnt, bool last, double *scal, Halo<S> hp) {
  if (!e->tma_ok || cnt > kTmaMaxCols) return false;
  const TmaPlan pl = tma_plan(e, cnt + 1, 0);
  const void *fn = last ? (const void *)k_lanczos_b_tma<S, true> : (const void *)k_lanczos_b_tma<S, false>;
  if (pl.nstages < 2 || !tma_attr(fn, e->smem_optin)) return false;
  const int gx = pl.ntile < e->nsm ? pl.ntile : e->nsm;
  if (last)
    k_lanczos_b_tma<S, true><<<dim3(gx, 1), kTmaThreads, pl.smem_bytes, e->stream>>>(e->g, pl, start, V, e->colstride, j, i0,
                                                                                   cnt, scal, e->rc, hp, e->status);
  else
    k_lanczos_b_tma<S, false><<<dim3(gx, 1), kTmaThreads, pl.smem_bytes, e->stream>>>(e->g, pl, start, V, e->colstride, j, i0,
                                                                                    cnt, scal, e->rc, hp, e->status);
  e->count(true);
  return true;
}

// ---------------------------------------------------------------- tensor-map pass A (nls_tmap.cuh)
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = []() -> EncodeTiledFn {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      return nullptr;
    }
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// 3-D map over a vector laid out like the state (`base` = first allocated element, ghost plane included for slabs):
// dims (doubles per row, ny, planes), box (doubles per (BX+2) sites, BY+2, 1), zero fill outside
bool tm_get(nls_engine *e, const void *base, int doubles_per_site, const CUtensorMap **out) {
  auto it = e->tmaps.find(base);
  if (it == e->tmaps.end()) {
    EncodeTiledFn enc = encode_tiled();
    if (!enc) return false;
    CUtensorMap tm;
    const Geom &g = e->g;
    const cuuint64_t planes = (cuuint64_t)g.nz_loc + (g.halo ? 2 : 0);
    const cuuint64_t dims[3] = {(cuuint64_t)g.nx * doubles_per_site, (cuuint64_t)g.ny, planes};
    const cuuint64_t strides[2] = {(cuuint64_t)g.nx * doubles_per_site * 8, (cuuint64_t)g.plane * doubles_per_site * 8};
    const cuuint32_t box[3] = {(cuuint32_t)(doubles_per_site == 2 ? TmRow<2>::PX * 2 : TmRow<1>::PX), (cuuint32_t)kTmPY, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return false;
    it = e->tmaps.emplace(base, tm).first;
  }
  *out = &it->second;
  return true;
}

// kernel for (mode, columns in the chunk); nullptr when there is none
constexpr int kTmMaxNC = 6; // two sites per thread: 2 * NC complex values in registers
template <class S, bool ANISO> const void *tm_kernel(int mode, int nc, bool pipe) {
  constexpr bool kHasPipe = sizeof(S) == 16; // the per-warp hand-back variant is built for complex states only
  if (pipe && !kHasPipe) pipe = false;
#define TMK(M, N)                                                                                       \
  if (mode == M && nc == N)                                                                             \
  return pipe ? (const void *)k_lanczos_a_tm<S, ANISO, N, M, kHasPipe> : (const void *)k_lanczos_a_tm<S, ANISO, N, M, false>
  TMK(TM_SELF, 1);
  TMK(TM_TAIL, 2); TMK(TM_TAIL, 3); TMK(TM_TAIL, 4); TMK(TM_TAIL, 5); TMK(TM_TAIL, 6);
  TMK(TM_BODY, 1); TMK(TM_BODY, 2); TMK(TM_BODY, 3); TMK(TM_BODY, 4); TMK(TM_BODY, 5); TMK(TM_BODY, 6);
#undef TMK
  return nullptr;
}
template <class S, bool ANISO> bool tm_attrs(int bytes) { // dynamic shared memory opt-in: not allowed later, inside a capture
  for (int mode = 0; mode < 3; ++mode)
    for (int nc = 1; nc <= kTmMaxNC; ++nc)
      for (int pipe = 0; pipe < 2; ++pipe)
        if (const void *fn = tm_kernel<S, ANISO>(mode, nc, pipe != 0))
          if (!tma_attr(fn, bytes)) return false;
  return true;
}

void tm_setup(nls_engine *e) {
  e->tm_ok = false;
  // default for complex 3D single-trajectory handles (measured with version 2: C5 61.7 -> 53.7 ms/step, C3 10.76 ->
  // 10.45, profiles/r02k_*); real-valued grids opt in with NLS_FLAG_TMAP / NLS_PASS_A_TMAP=1; NLS_PASS_A_TMAP=0 or
  // NLS_NO_TMAP=1 force the flat pass
  const char *env = getenv("NLS_PASS_A_TMAP");
  const bool want = env ? atoi(env) != 0 : ((e->d.flags & NLS_FLAG_TMAP) || e->cplx);
  if (!want || (e->d.flags & NLS_FLAG_NO_TMAP) || getenv("NLS_NO_TMAP")) return;
  const Geom &g = e->g;
  if (!g.dim3 || e->batch != 1 || g.nx < 4 || g.ny < 4) return;
  if (!e->cplx && (g.nx % 2)) return; // row pitch must be a multiple of 16 bytes
  if (!encode_tiled()) return;
  const bool an = e->d.laplacian == NLS_LAP_ANISO;
  const int bytes = (int)tm_smem_bytes(e->W, an, kTmRPipe); // opt-in for the larger of the two ring sizes
  const bool ok = e->cplx ? (an ? tm_attrs<double2, true>(bytes) : tm_attrs<double2, false>(bytes))
                          : (an ? tm_attrs<double, true>(bytes) : tm_attrs<double, false>(bytes));
  e->tm_ok = ok;
}

// First pass-A chunk of iteration j over the columns c0..j through the tensor-map kernel.  Chooses the chunk
// (*cnt columns from c0) so that the role of every column is a compile-time fact (nls_tmap.cuh): all of c0..j when
// they fit (TM_TAIL; TM_SELF for one column), else the leading columns below j-1 (TM_BODY).
template <class S, bool ANISO>
bool launch_a_tm(nls_engine *e, const S *start, S *V, int j, int c0, int cap, int *cnt, double *scal, bool raw = false) {
  if (!e->tm_ok) return false;
  if (cap > kTmMaxNC) cap = kTmMaxNC;
  const int total = j + 1 - c0;
  int mode, nc;
  if (total == 1) { mode = TM_SELF; nc = 1; }
  else if (total <= cap) { mode = TM_TAIL; nc = total; }
  else {
    // The split between this chunk and the dots-only chunk that follows moves no bytes (the second chunk re-reads w
    // either way), so the body chunk is kept at the width that still fits 128 registers without spills
    // (6 columns: 184-208 bytes of stack per thread, 4: none); NLS_TM_BODY overrides for tuning.
    static const int body = []() { const char *b = getenv("NLS_TM_BODY"); const int v = b ? atoi(b) : 4; return v >= 1 && v <= kTmMaxNC ? v : 4; }();
    if (cap > body) cap = body;
    mode = TM_BODY;
    nc = j - 1 - c0 < cap ? j - 1 - c0 : cap;
  }
  // ring hand-back style and L2 prefetch lead of the column values: tuning knobs (NLS_TM_PIPE, NLS_TM_PREFETCH)
  static const int pipe_knob = []() { const char *v = getenv("NLS_TM_PIPE"); return v ? atoi(v) != 0 : kTmDefaultPipe; }();
  static const int tpf = []() { const char *v = getenv("NLS_TM_PREFETCH"); const int n = v ? atoi(v) : kTmDefaultPrefetch; return n >= 0 && n <= 8 ? n : 0; }();
  const bool pipe = pipe_knob && sizeof(S) == 16;
  const void *fn = nc >= 1 ? tm_kernel<S, ANISO>(mode, nc, pipe) : nullptr;
  if (!fn) return false;
  Geom g = e->g;
  g.raw = raw ? 1 : 0;
  const S *cj = j == 0 ? start : V + (size_t)(j - 1) * e->colstride; // base of column j (halo included)
  const CUtensorMap *tmx = nullptr, *tmc = nullptr;
  if (!tm_get(e, cj, (int)(sizeof(S) / 8), &tmx)) return false;
  if (ANISO) { if (!tm_get(e, e->cf, 1, &tmc)) return false; }
  else tmc = tmx;
  const dim3 gr = grid_for(e, fn);
  TmGeom tg;
  tg.ntx = (g.nx + kTmBX - 1) / kTmBX;
  tg.nty = (g.ny + kTmBY - 1) / kTmBY;
  const long long tiles = (long long)tg.ntx * tg.nty;
  long long nzc = (3LL * gr.x + tiles - 1) / tiles;
  if (nzc < 1) nzc = 1;
  int zc = (int)((g.nz_loc + nzc - 1) / nzc);
  if (zc < 8) zc = 8;
  if (zc > g.nz_loc) zc = g.nz_loc;
  tg.zc = zc;
  tg.nzc = (g.nz_loc + zc - 1) / zc;
  tg.units = tiles * tg.nzc;
  tg.zoff = g.halo ? 1 : 0;
  tg.pf = tpf;
  CUtensorMap mx = *tmx, mc = *tmc;
  size_t colstride = e->colstride;
  int i0 = c0;
  const double *cf = e->cf;
  RedCtx rc = e->rc;
  int *status = e->status;
  void *args[] = {&mx, &mc, &g, &tg, &start, &V, &colstride, &j, &i0, &cf, &scal, &rc, &status};
  if (cudaLaunchKernel(fn, dim3(gr.x, 1), dim3(kThreads), args, tm_smem_bytes((int)(sizeof(S) / 8), ANISO, pipe ? kTmRPipe : kTmR), e->stream) != cudaSuccess)
    return false;
  e->count();
  e->tm_launches += 1;
  *cnt = nc;
  return true;
}

template <class S, bool ANISO>
void lanczos_t(nls_engine *e, const S *start, double *scal) {
  const int m = e->m;
  const dim3 gr = e->grid();
  S *V = (S *)e->V;
  constexpr int kNC = 12; // columns per launch
  // NLS_REORTH_GRAM: one narrow stencil pass S(j) (raw r_j = L column_j, true dots with columns j-1 and j only) and one
  // update pass U(j+1) over all columns whose older projections come from Gram entries (lanczos_b_body<GDR>): j + 6
  // vector passes per iteration instead of 2j + 5.  All columns must fit one closing chunk (m <= kNC + 1).
  const bool gdr = m - 1 <= kNC && (e->d.reorth == NLS_REORTH_GRAM || (e->d.reorth == NLS_REORTH_FULL && e->gdr_auto));
  for (int j = 0; gdr && j < m - 1; ++j) {
    const int c0 = j > 0 ? j - 1 : 0, total = j + 1 - c0;
    int cnt = 0;
    if (!launch_a_tm<S, ANISO>(e, start, V, j, c0, 2, &cnt, scal, true)) {
      Geom graw = e->g;
      graw.raw = 1;
      LAUNCH(e, (k_lanczos_a<S, ANISO, 2, true>), gr, graw, start, V, e->colstride, j, c0, total, e->cf, scal, e->rc);
    }
    const Halo<S> hp = halo_for<S>(e, V + (size_t)j * e->colstride);
#define LG(NC) LAUNCH(e, (k_lanczos_b_gdr<S, NC>), gr, e->g, start, V, e->colstride, j, 0, j + 1, scal, e->rc, hp)
    if (j + 1 <= 2) LG(2);
    else if (j + 1 <= 4) LG(4);
    else if (j + 1 <= 6) LG(6);
    else if (j + 1 <= 8) LG(8);
    else if (j + 1 <= 10) LG(10);
    else LG(12);
#undef LG
  }
  if (gdr) return;
  const int reorth = (e->d.reorth == NLS_REORTH_GRAM || e->d.reorth == NLS_REORTH_BLOCK) ? NLS_REORTH_FULL : e->d.reorth;
  for (int j = 0; j < m - 1; ++j) {
    // columns taking part in the orthogonalisation: all of 0..j (reference), or only j (3-term)
    const int c0 = reorth != NLS_REORTH_FULL ? j : 0;
    // pass A chunks: the stencil-carrying first chunk of a complex sweep takes at most kFirst columns;
    // beyond that the 128-register budget of 2 CTAs/SM costs more than re-reading w once in a second,
    // dots-only chunk.  Measured on B200 (profiles/r01f_firstcap.txt): div(c grad) stencil 6 (C3 12.0 ->
    // 10.8 ms/step), uniform stencil 10 (C4/C5 lose 3% with 6).  NLS_FIRST_CAP overrides for tuning.
    const int kFirst = sizeof(S) == 16 ? (e->first_cap >= 2 && e->first_cap <= kNC ? e->first_cap : ((ANISO || e->tm_ok) ? 6 : 10)) : kNC;
    for (int i0 = c0, cnt = 0; i0 <= j; i0 += cnt) {
      const bool first = i0 == c0;
      const int cap = first ? kFirst : kNC;
      cnt = j + 1 - i0 < cap ? j + 1 - i0 : cap;
      if (first && i0 == 0 && cnt == j + 1 && launch_a_tma<S, ANISO>(e, start, V, j, scal)) continue;
      if (first && !(e->g.dim3 && (e->d.flags & NLS_FLAG_MARCH)) && launch_a_tm<S, ANISO>(e, start, V, j, c0, cap, &cnt, scal)) continue;
#define LA(NC)                                                                                           \
  do {                                                                                                   \
    if (first && i0 == 0 && e->g.dim3 && (e->d.flags & NLS_FLAG_MARCH)) {                            \
      const dim3 g3 = grid_for(e, (const void *)k_lanczos_a3<S, ANISO, NC>);                             \
      const March mc = make_march(e->g, 8 * (int)g3.x);                                                  \
      k_lanczos_a3<S, ANISO, NC><<<g3, kThreads, 0, e->stream>>>(e->g, mc, start, V, e->colstride, j, cnt, e->cf, scal, e->rc); \
      e->count();                                                                                        \
    } else if (first) LAUNCH(e, (k_lanczos_a<S, ANISO, NC, true>), gr, e->g, start, V, e->colstride, j, i0, cnt, e->cf, scal, e->rc);  \
    else LAUNCH(e, (k_lanczos_a<S, ANISO, NC, false>), gr, e->g, start, V, e->colstride, j, i0, cnt, e->cf, scal, e->rc);       \
  } while (0)
      if (cnt <= 2) LA(2);
      else if (cnt <= 4) LA(4);
      else if (cnt <= 6) LA(6);
      else if (cnt <= 8) LA(8);
      else if (cnt <= 10) LA(10);
      else LA(12);
#undef LA
    }
    if (reorth == NLS_REORTH_MGS) { // sequential sweep of the reference, one dot per launch (k_lanczos_mgs)
      const Halo<S> none{nullptr, nullptr};
      for (int i = 0; i <= j; ++i) // subtract the previous projection (first: the alpha term), dot with column i
        LAUNCH(e, (k_lanczos_mgs<S, false>), gr, e->g, start, V, e->colstride, j, i == 0 ? j : i - 1, i, i == 0 ? 1 : 0, scal, e->rc, none);
      LAUNCH(e, (k_lanczos_mgs<S, true>), gr, e->g, start, V, e->colstride, j, j, 0, 0, scal, e->rc,
             halo_for<S>(e, V + (size_t)j * e->colstride));
      continue;
    }
    // Gram entries of column j against the columns the LAST chunk of the previous pass B did not hold
    if (c0 == 0 && j >= 1)
      for (int i0 = 0; i0 < ((j - 1) / kNC) * kNC; i0 += kNC)
        LAUNCH(e, (k_lanczos_g<S, kNC>), gr, e->g, start, (const S *)V, e->colstride, j, i0, kNC, scal, e->rc);
    for (int i0 = c0; i0 <= j; i0 += kNC) {
      const int cnt = j + 1 - i0 < kNC ? j + 1 - i0 : kNC;
      const bool last = i0 + kNC > j;
      const Halo<S> hp = last ? halo_for<S>(e, V + (size_t)j * e->colstride) : Halo<S>{nullptr, nullptr};
      if (launch_b_tma<S>(e, start, V, j, i0, cnt, last, scal, hp)) continue;
#define LB(NC)                                                                                           \
  do {                                                                                                   \
    if (last) LAUNCH(e, (k_lanczos_b<S, NC, true>), gr, e->g, start, V, e->colstride, j, i0, cnt, scal, e->rc, hp);  \
    else LAUNCH(e, (k_lanczos_b<S, NC, false>), gr, e->g, start, V, e->colstride, j, i0, cnt, scal, e->rc, hp);      \
  } while (0)
      if (cnt <= 2) LB(2);
      else if (cnt <= 4) LB(4);
      else if (cnt <= 6) LB(6);
      else if (cnt <= 8) LB(8);
      else if (cnt <= 10) LB(10);
      else LB(12);
#undef LB
    }
  }
}

template <class S>
void lanczos(nls_engine *e, const S *start, double *scal) {
  if (e->d.laplacian == NLS_LAP_ANISO) lanczos_t<S, true>(e, start, scal);
  else lanczos_t<S, false>(e, start, scal);
}

void tridiag(nls_engine *e, const double *scal, int nfn, int f0, double t0r, double t0i, int f1,
             double t1r, double t1i) {
  TridiagArgs a;
  a.m = e->m;
  a.W = e->W;
  a.nfn = nfn;
  a.fn[0] = f0; a.tr[0] = t0r; a.ti[0] = t0i;
  a.fn[1] = f1; a.tr[1] = t1r; a.ti[1] = t1i;
  k_tridiag<<<e->batch, 32, 0, e->stream>>>(scal, e->coef, a, e->status);
  e->count();
}

// ---------------------------------------------------------------- one time step
template <int EQ>
void nlse_step_bufs(nls_engine *e, double dt, const double2 *uin, double2 *uout, const double2 *scur, double2 *snext) {
  lanczos<double2>(e, scur, e->scalA);
  tridiag(e, e->scalA, 1, NLS_FN_EXP, 0.0, dt, 0, 0, 0); // expm_multiply(L, rho, tau = i dt)
  NlseParams p{0.5 * dt, e->d.s1, e->d.s2, e->d.kappa};
  LAUNCH(e, (k_nlse_finish<EQ>), e->grid(), e->g, scur, (const double2 *)e->V, e->colstride,
         e->coef, e->m, uin, e->mf, uout, snext, p, e->scalA, e->rc, halo_for<double2>(e, snext));
}

template <int EQ>
void nlse_step_t(nls_engine *e, double dt) {
  // start-vector ping-pong (column 0 of the basis is read at clamped sites while the next rho is written)
  nlse_step_bufs<EQ>(e, dt, (const double2 *)e->u[e->phase], (double2 *)e->u[1 - e->phase],
                     (const double2 *)e->sv(e->phase), (double2 *)e->sv(1 - e->phase));
}

// sEWI (NLS_INT_SEWI): three rotating state buffers u = buf[p], u- = buf[(p+2)%3], u+ -> buf[(p+1)%3].
// First step after a state upload: one SS2 step with u- = u0 (nlse_cubic_gautschi_driver.cpp:122-125).
template <int EQ>
void sewi_first_step_t(nls_engine *e, double dt) {
  NlseParams p{0.5 * dt, e->d.s1, e->d.s2, e->d.kappa};
  const double2 *u = (const double2 *)e->u[e->phase];
  LAUNCH(e, (k_nlse_prephase<EQ>), e->grid(), e->g, u, e->mf, (double2 *)e->w0, p, e->scalA, e->rc,
         halo_for<double2>(e, e->w0));
  nlse_step_bufs<EQ>(e, dt, u, (double2 *)e->u[(e->phase + 1) % 3], (const double2 *)e->w0, (double2 *)e->wbuf[0]);
}

// u+ = e^{sg 2 tau L} u- - sg 2 tau e^{sg tau L} sinc(dt L) B(u)   (sg = -1: NLS_FLAG_SEWI_HOST2D)
// nlse_cubic_gautschi_solver_3d.hpp:8-35 / nlse_cubic_gautschi_solver.hpp:21-39 / nlse_dev.hpp:205-238
template <int EQ>
void sewi_step_t(nls_engine *e, double dt) {
  const double sg = (e->d.flags & NLS_FLAG_SEWI_HOST2D) ? -1.0 : 1.0;
  const double2 *u = (const double2 *)e->u[e->phase];
  const double2 *uprev = (const double2 *)e->u[(e->phase + 2) % 3];
  double2 *unew = (double2 *)e->u[(e->phase + 1) % 3];
  double2 *B = (double2 *)e->w0, *phi = (double2 *)e->wbuf[0], *a = (double2 *)e->acc;
  const double2 *V = (const double2 *)e->V;
  NlseParams p{0.5 * dt, e->d.s1, e->d.s2, e->d.kappa};
  LAUNCH(e, (k_sewi_B<EQ>), e->grid(), e->g, u, e->mf, B, p, e->scalA, e->rc, halo_for<double2>(e, B));
  lanczos<double2>(e, B, e->scalA);
  tridiag(e, e->scalA, 1, NLS_FN_SINC, dt, 0.0, 0, 0, 0); // sincm_multiply(L, B, dt)
  LAUNCH(e, (k_combine_norm<double2>), e->grid(), e->g, (const double2 *)B, V, e->colstride, e->coef, e->m, phi,
         e->scalA, e->rc, halo_for<double2>(e, phi));
  lanczos<double2>(e, phi, e->scalA);
  tridiag(e, e->scalA, 1, NLS_FN_EXP, 0.0, sg * dt, 0, 0, 0); // expm_multiply(L, phi, sg tau)
  LAUNCH(e, (k_combine<double2>), e->grid(), e->g, (const double2 *)phi, V, e->colstride, e->coef, 0, e->m, a);
  halo_exchange<double2>(e, uprev); // slabs: u- becomes a stencil input
  LAUNCH(e, (k_norm2<double2>), e->grid(), e->g, uprev, e->scalA + kOffNrm2, kScalStride, e->rc);
  lanczos<double2>(e, uprev, e->scalA);
  tridiag(e, e->scalA, 1, NLS_FN_EXP, 0.0, sg * 2.0 * dt, 0, 0, 0); // expm_multiply(L, u-, sg 2 tau)
  LAUNCH(e, k_sewi_final, e->grid(), e->g, uprev, V, e->colstride, e->coef, e->m, (const double2 *)a, unew,
         -sg * 2.0 * dt);
}

template <int EQ>
void gautschi_step_t(nls_engine *e, double dt) {
  double *u = (double *)e->u[0], *up = (double *)e->u[1];
  lanczos<double>(e, u, e->scalA);
  const int filt = EQ == NLS_EQ_SGE ? NLS_FN_MOD_COSINE : NLS_FN_ID_SQRT;
  tridiag(e, e->scalA, 2, filt, dt, 0.0, NLS_FN_COS_SQRT, dt, 0.0);
  LAUNCH(e, (k_gautschi_a<EQ>), e->grid(), e->g, (const double *)e->V, e->colstride,
         e->coef, e->m, u, up, e->mf, (double *)e->w0, (double *)e->acc, e->scalB,
         e->rc, halo_for<double>(e, e->w0));
  lanczos<double>(e, (const double *)e->w0, e->scalB);
  tridiag(e, e->scalB, 1, NLS_FN_SINC2_HALF, dt, 0.0, 0, 0, 0);
  LAUNCH(e, (k_gautschi_b<0>), e->grid(), e->g, (const double *)e->w0, (const double *)e->V, e->colstride,
         e->coef, e->m, (const double *)e->acc, u, dt * dt, e->scalA, e->rc,
         halo_for<double>(e, u));
}

template <int EQ>
void sv_step_t(nls_engine *e, double dt) {
  // rotating buffers: u = buf[phase], u_past = buf[(phase+2)%3], new = buf[(phase+1)%3]
  const double *u = (const double *)e->u[e->phase];
  const double *up = (const double *)e->u[(e->phase + 2) % 3];
  double *un = (double *)e->u[(e->phase + 1) % 3];
  if (e->d.laplacian == NLS_LAP_ANISO)
    LAUNCH(e, (k_sv<EQ, true>), e->grid(), e->g, u, up, e->mf, e->cf, un, dt * dt, e->rc, halo_for<double>(e, un),
           e->slab_scratch);
  else
    LAUNCH(e, (k_sv<EQ, false>), e->grid(), e->g, u, up, e->mf, e->cf, un, dt * dt, e->rc, halo_for<double>(e, un),
           e->slab_scratch);
}

void issue_step(nls_engine *e, double dt) {
  const int eq = e->d.equation;
  if (e->cplx && e->d.integrator == NLS_INT_SEWI) {
    if (!e->sewi_started) {
      if (eq == NLS_EQ_NLSE_CUBIC) sewi_first_step_t<NLS_EQ_NLSE_CUBIC>(e, dt);
      else if (eq == NLS_EQ_NLSE_CQ) sewi_first_step_t<NLS_EQ_NLSE_CQ>(e, dt);
      else sewi_first_step_t<NLS_EQ_NLSE_SAT>(e, dt);
    } else {
      if (eq == NLS_EQ_NLSE_CUBIC) sewi_step_t<NLS_EQ_NLSE_CUBIC>(e, dt);
      else if (eq == NLS_EQ_NLSE_CQ) sewi_step_t<NLS_EQ_NLSE_CQ>(e, dt);
      else sewi_step_t<NLS_EQ_NLSE_SAT>(e, dt);
    }
    return;
  }
  if (e->cplx) {
    if (eq == NLS_EQ_NLSE_CUBIC) nlse_step_t<NLS_EQ_NLSE_CUBIC>(e, dt);
    else if (eq == NLS_EQ_NLSE_CQ) nlse_step_t<NLS_EQ_NLSE_CQ>(e, dt);
    else nlse_step_t<NLS_EQ_NLSE_SAT>(e, dt);
    return;
  }
#define REAL_DISPATCH(fn)                                                        \
  switch (eq) {                                                                  \
  case NLS_EQ_SGE: fn<NLS_EQ_SGE>(e, dt); break;                                 \
  case NLS_EQ_KGE: fn<NLS_EQ_KGE>(e, dt); break;                                 \
  case NLS_EQ_PHI4: fn<NLS_EQ_PHI4>(e, dt); break;                               \
  case NLS_EQ_SGE_DOUBLE: fn<NLS_EQ_SGE_DOUBLE>(e, dt); break;                   \
  default: fn<NLS_EQ_SGE_SINH>(e, dt); break;                                    \
  }
  if (e->d.integrator == NLS_INT_SV) { REAL_DISPATCH(sv_step_t) }
  else { REAL_DISPATCH(gautschi_step_t) }
#undef REAL_DISPATCH
}

// current-state buffers (real equations)
double *cur_u(nls_engine *e) {
  return (double *)(e->d.integrator == NLS_INT_SV ? e->u[e->phase] : e->u[0]);
}
double *cur_upast(nls_engine *e) {
  return (double *)(e->d.integrator == NLS_INT_SV ? e->u[(e->phase + 2) % 3] : e->u[1]);
}

// make the start vector of the next step valid (first step, or after a state/dt change)
int ensure_start(nls_engine *e, double dt) {
  if (e->d.integrator == NLS_INT_SV || e->d.integrator == NLS_INT_SEWI) return NLS_OK;
  if (e->start_valid && (!e->cplx || e->start_dt == dt)) return NLS_OK;
  if (e->cplx) {
    NlseParams p{0.5 * dt, e->d.s1, e->d.s2, e->d.kappa};
    const double2 *u = (const double2 *)e->u[e->phase];
    switch (e->d.equation) {
    case NLS_EQ_NLSE_CUBIC:
      LAUNCH(e, (k_nlse_prephase<NLS_EQ_NLSE_CUBIC>), e->grid(), e->g, u, e->mf, (double2 *)e->sv(e->phase), p, e->scalA, e->rc,
             halo_for<double2>(e, e->sv(e->phase)));
      break;
    case NLS_EQ_NLSE_CQ:
      LAUNCH(e, (k_nlse_prephase<NLS_EQ_NLSE_CQ>), e->grid(), e->g, u, e->mf, (double2 *)e->sv(e->phase), p, e->scalA, e->rc,
             halo_for<double2>(e, e->sv(e->phase)));
      break;
    default:
      LAUNCH(e, (k_nlse_prephase<NLS_EQ_NLSE_SAT>), e->grid(), e->g, u, e->mf, (double2 *)e->sv(e->phase), p, e->scalA, e->rc,
             halo_for<double2>(e, e->sv(e->phase)));
    }
  } else {
    LAUNCH(e, (k_norm2<double>), e->grid(), e->g, (const double *)e->u[0], e->scalA + kOffNrm2,
           kScalStride, e->rc);
  }
  CU(cudaGetLastError());
  e->start_valid = true;
  e->start_dt = dt;
  return NLS_OK;
}

int run_steps_pk(nls_engine *e, double dt, uint32_t nsteps);
int run_steps_sv2(nls_engine *e, double dt, uint32_t nsteps);

int run_steps(nls_engine *e, double dt, uint32_t nsteps) {
  if (nsteps == 0) return NLS_OK;
  if (e->sv2_ok) return run_steps_sv2(e, dt, nsteps);
  if (e->pk_ok) return run_steps_pk(e, dt, nsteps);
  int rc = ensure_start(e, dt);
  if (rc) return rc;
  const bool use_graph = !(e->d.flags & NLS_FLAG_NO_GRAPH);
  for (uint32_t s = 0; s < nsteps; ++s) {
    const int p = e->phase;
    if (e->d.integrator == NLS_INT_SEWI && !e->sewi_started) { // one-off SS2 start-up step, not worth a graph
      issue_step(e, dt);
      CU(cudaGetLastError());
      e->sewi_started = true;
      e->phase = (p + 1) % e->nphase;
      continue;
    }
    if (use_graph) {
      if (e->gexec[p] && e->gdt[p] != dt) {
        cudaGraphExecDestroy(e->gexec[p]);
        e->gexec[p] = nullptr;
      }
      if (!e->gexec[p]) {
        cudaGraph_t graph;
        CU(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
        e->capturing = true;
        e->cap_count = 0;
        e->cap_tma = 0;
        issue_step(e, dt);
        e->capturing = false;
        CU(cudaStreamEndCapture(e->stream, &graph));
        CU(cudaGraphInstantiate(&e->gexec[p], graph, 0));
        cudaGraphDestroy(graph);
        e->gdt[p] = dt;
        e->glaunch[p] = e->cap_count;
        e->gtma[p] = e->cap_tma;
      }
      CU(cudaGraphLaunch(e->gexec[p], e->stream));
      e->launches += e->glaunch[p];
      e->tma_launches += e->gtma[p];
    } else {
      issue_step(e, dt);
      CU(cudaGetLastError());
    }
    e->phase = (p + 1) % e->nphase;
  }
  return NLS_OK;
}

// ---- persistent path: one cooperative launch advances nsteps steps
void pk_setup(nls_engine *e) {
  const nls_desc &d = e->d;
  e->pk_ok = false;
  if (d.flags & NLS_FLAG_NO_PERSISTENT) return;
  const bool sv = d.integrator == NLS_INT_SV;
  if (sv && !(d.flags & NLS_FLAG_PK_SV)) return; // measured: one barrier per step costs more than a k_sv launch
  if (e->batch != 1 || e->g.halo != 0 || (!sv && d.reorth != NLS_REORTH_FULL)) return;
  if (d.integrator != NLS_INT_SS2 && d.integrator != NLS_INT_GAUTSCHI && !sv) return;
  int dev = d.device, nsm = 0, smem_max = 0, coop = 0;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
  if (!coop || nsm <= 0 || nsm > kPkPartStride) return;
  long long csz = (e->g.n + nsm - 1) / nsm;
  // Fewer, larger CTAs on small grids were tried and lose: a grid-wide round costs 2.2 us + 0.23 us per reduced value at
  // 148 CTAs (profiles/r02d_allreduce_microbench.log) and gets cheaper with fewer CTAs, but the site loops between the
  // rounds are L2-latency chains that grow faster: C1 0.100 ms/step at 148 CTAs, 0.108 at 74, 0.125 at 37, 0.165 at 16
  // (profiles/r02o_pk_min_csz_sweep.txt).  NLS_PK_MIN_CSZ (minimum sites per CTA) stays as the tuning knob.
  static const long long min_csz = []() { const char *v = getenv("NLS_PK_MIN_CSZ"); const long long n = v ? atoll(v) : kPkMinSites; return n >= 64 ? n : 64; }();
  if (csz < min_csz) csz = min_csz;
  const int ncta = (int)((e->g.n + csz - 1) / csz);
  const size_t smem = pk_smem_bytes(e->esz, csz, e->g.nx, e->m, sv ? 2 * e->g.nx : 0);
  if (smem > (size_t)smem_max) return;
  const bool an = d.laplacian == NLS_LAP_ANISO;
  const void *fn = e->cplx ? (an ? (const void *)k_pk_nlse<true> : (const void *)k_pk_nlse<false>)
                   : sv    ? (an ? (const void *)k_pk_sv<true> : (const void *)k_pk_sv<false>)
                           : (an ? (const void *)k_pk_gautschi<true> : (const void *)k_pk_gautschi<false>);
  if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    cudaGetLastError();
    return;
  }
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, kPkThreads, smem) != cudaSuccess || occ < 1) {
    cudaGetLastError();
    return;
  }
  if (cudaMalloc((void **)&e->pk_part, (size_t)2 * kPkMaxK * kPkPartStride * 8) != cudaSuccess) return;
  cudaMemset(e->pk_part, 0, (size_t)2 * kPkMaxK * kPkPartStride * 8);
  if (cudaMalloc((void **)&e->pk_bar, 16) != cudaSuccess) return;
  e->pk_ncta = ncta;
  e->pk_csz = csz;
  e->pk_smem = smem;
  e->pk_ok = true;
}

int run_steps_pk_once(nls_engine *e, double dt, uint32_t nsteps);
// The grid-barrier counter is 32-bit and only grows within a launch (ncta arrivals per barrier, ~40 barriers
// per step): split long runs so it can never wrap (148 CTAs x 40 barriers x 65536 steps = 3.9e8 < 2^32).
int run_steps_pk(nls_engine *e, double dt, uint32_t nsteps) {
  constexpr uint32_t kMaxStepsPerLaunch = 65536u / 2;
  while (nsteps) {
    const uint32_t k = nsteps < kMaxStepsPerLaunch ? nsteps : kMaxStepsPerLaunch;
    const int rc = run_steps_pk_once(e, dt, k);
    if (rc) return rc;
    nsteps -= k;
  }
  return NLS_OK;
}

int run_steps_pk_once(nls_engine *e, double dt, uint32_t nsteps) {
  PkArgs a;
  a.g = e->g;
  a.m = e->m; a.eq = e->d.equation; a.integ = e->d.integrator;
  a.aniso = e->d.laplacian == NLS_LAP_ANISO;
  a.debug_fault = (e->d.flags & NLS_FLAG_DEBUG_PK_FAULT) ? 1 : 0;
  a.nsteps = (int)nsteps; a.ncta = e->pk_ncta; a.csz = e->pk_csz;
  a.dt = dt; a.s1 = e->d.s1; a.s2 = e->d.s2; a.kappa = e->d.kappa;
  a.V = e->V; a.colstride = e->colstride;
  a.acc = e->acc; a.mf = e->mf; a.cf = e->cf;
  a.part = e->pk_part; a.bar = e->pk_bar; a.status = e->status;
  CU(cudaMemsetAsync(e->pk_bar, 0, 16, e->stream));
  const bool an = a.aniso;
  if (e->cplx) {
    a.u0 = e->u[e->phase]; a.u1 = e->u[1 - e->phase];
    double2 *sv0 = (double2 *)e->w0, *sv1 = (double2 *)e->wbuf[0];
    void *args[] = {&a, &sv0, &sv1};
    const void *fn = an ? (const void *)k_pk_nlse<true> : (const void *)k_pk_nlse<false>;
    CU(cudaLaunchCooperativeKernel(fn, dim3(a.ncta), dim3(kPkThreads), args, e->pk_smem, e->stream));
    e->phase = (e->phase + (int)(nsteps & 1u)) & 1;
  } else if (e->d.integrator == NLS_INT_SV) {
    double *b0 = (double *)e->u[e->phase], *b1 = (double *)e->u[(e->phase + 1) % 3], *b2 = (double *)e->u[(e->phase + 2) % 3];
    void *args[] = {&a, &b0, &b1, &b2};
    const void *fn = an ? (const void *)k_pk_sv<true> : (const void *)k_pk_sv<false>;
    CU(cudaLaunchCooperativeKernel(fn, dim3(a.ncta), dim3(kPkThreads), args, e->pk_smem, e->stream));
    e->phase = (int)((e->phase + nsteps) % 3u);
  } else {
    a.u0 = e->u[0]; a.u1 = e->u[1];
    double *gb = (double *)e->w0;
    void *args[] = {&a, &gb};
    const void *fn = an ? (const void *)k_pk_gautschi<true> : (const void *)k_pk_gautschi<false>;
    CU(cudaLaunchCooperativeKernel(fn, dim3(a.ncta), dim3(kPkThreads), args, e->pk_smem, e->stream));
  }
  e->launches += 1;
  e->start_valid = false; // the multi-kernel path's start vector / norms are stale now
  return NLS_OK;
}

// ---- temporally blocked Stoermer-Verlet (2D, uniform Laplacian, one trajectory): default for every grid whose
// tiles fit in shared memory; NLS_FLAG_PK_SV selects the older 1-D-chunk kernel instead, NLS_FLAG_NO_PERSISTENT the
// one-kernel-per-step graph
template <int EQ> const void *sv2_fn() { return (const void *)k_pk_sv2<EQ>; }
const void *sv2_kernel(int eq) {
  switch (eq) {
  case NLS_EQ_SGE: return sv2_fn<NLS_EQ_SGE>();
  case NLS_EQ_KGE: return sv2_fn<NLS_EQ_KGE>();
  case NLS_EQ_PHI4: return sv2_fn<NLS_EQ_PHI4>();
  case NLS_EQ_SGE_DOUBLE: return sv2_fn<NLS_EQ_SGE_DOUBLE>();
  default: return sv2_fn<NLS_EQ_SGE_SINH>();
  }
}

void sv2_setup(nls_engine *e) {
  const nls_desc &d = e->d;
  e->sv2_ok = false;
  if (d.integrator != NLS_INT_SV || (d.flags & (NLS_FLAG_NO_PERSISTENT | NLS_FLAG_PK_SV))) return;
  if (e->batch != 1 || e->g.dim3 || e->slab || d.laplacian != NLS_LAP_UNIFORM) return;
  int dev = d.device, nsm = 0, smem_max = 0, coop = 0;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
  if (!coop || nsm <= 0) return;
  Sv2Plan pl;
  if (!sv2_plan(e->g.nx, e->g.ny, nsm, (size_t)smem_max - 1024, pl)) return;
  if (const char *ms = getenv("NLS_SV2_MAX_S")) { // tuning knob, not API: cap the temporal block depth
    const int cap = atoi(ms);
    if (cap >= 1 && pl.s > cap) { pl.s = cap; pl.smem = sv2_smem_bytes(pl.w, pl.h, pl.s); }
  }
  const void *fn = sv2_kernel(d.equation);
  if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem) != cudaSuccess) {
    cudaGetLastError();
    return;
  }
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, kPkThreads, pl.smem) != cudaSuccess || occ < 1) {
    cudaGetLastError();
    return;
  }
  if (pl.px * pl.py > nsm * occ) return;
  if (cudaMalloc((void **)&e->sv2_prog, (size_t)(pl.px * pl.py + 2) * 8) != cudaSuccess) return;
  e->sv2 = pl;
  e->sv2_ok = true;
}

int run_steps_sv2(nls_engine *e, double dt, uint32_t nsteps) {
  const Sv2Plan &pl = e->sv2;
  const int ncta = pl.px * pl.py;
  const int p = e->phase, pp = (p + 2) % 3, pn = (p + 1) % 3;
  Sv2Args a;
  a.nx = e->g.nx; a.ny = e->g.ny; a.nsteps = (int)nsteps; a.pl = pl;
  a.sc = e->g.sc; a.tau2 = dt * dt; a.mf = e->mf;
  // buffer sets: [0] = current (u, u_past); [1], [2] = ping-pong outputs (spare state buffer + Krylov work vectors)
  void **slot_u[3] = {&e->u[p], &e->u[pn], &e->wbuf[1]};
  void **slot_p[3] = {&e->u[pp], &e->wbuf[0], &e->w0};
  for (int q = 0; q < 3; ++q) { a.su[q] = (double *)*slot_u[q]; a.sp[q] = (double *)*slot_p[q]; }
  a.prog = e->sv2_prog;
  a.abort_flag = (unsigned *)(e->sv2_prog + ncta);
  a.status = e->status;
  a.debug_fault = (e->d.flags & NLS_FLAG_DEBUG_PK_FAULT) ? 1 : 0;
  CU(cudaMemsetAsync(e->sv2_prog, 0, (size_t)(ncta + 2) * 8, e->stream));
  void *args[] = {&a};
  CU(cudaLaunchCooperativeKernel(sv2_kernel(e->d.equation), dim3(ncta), dim3(kPkThreads), args, pl.smem, e->stream));
  e->launches += 1;
  // the last block (index nb-1) stored the final state in set 1 + (nb-1) % 2: make it the current state by
  // exchanging buffer roles (pointers only)
  const int nb = ((int)nsteps + pl.s - 1) / pl.s;
  const int fs = 1 + ((nb - 1) & 1);
  std::swap(*slot_u[0], *slot_u[fs]);
  std::swap(*slot_p[0], *slot_p[fs]);
  e->start_valid = false;
  return NLS_OK;
}

// host <-> device copies of [batch][n] host arrays into [batch][stride] (+halo) device arrays
int copy_in(nls_engine *e, void *dev, const void *host, size_t esz, int nb, bool from_dev) {
  const cudaMemcpyKind k = from_dev ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  if (e->g.stride == e->g.n) {
    CU(cudaMemcpyAsync(dev, host, (size_t)nb * e->g.n * esz, k, e->stream));
    return NLS_OK;
  }
  for (int b = 0; b < nb; ++b)
    CU(cudaMemcpyAsync((char *)dev + ((size_t)b * e->g.stride + e->g.halo) * esz,
                       (const char *)host + (size_t)b * e->g.n * esz, e->g.n * esz, k, e->stream));
  return NLS_OK;
}
int copy_out(nls_engine *e, void *host, const void *dev, size_t esz, int nb, bool to_dev) {
  const cudaMemcpyKind k = to_dev ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
  if (e->g.stride == e->g.n) {
    CU(cudaMemcpyAsync(host, dev, (size_t)nb * e->g.n * esz, k, e->stream));
  } else {
    for (int b = 0; b < nb; ++b)
      CU(cudaMemcpyAsync((char *)host + (size_t)b * e->g.n * esz,
                         (const char *)dev + ((size_t)b * e->g.stride + e->g.halo) * esz,
                         e->g.n * esz, k, e->stream));
  }
  CU(cudaStreamSynchronize(e->stream));
  return NLS_OK;
}

// Device-side failures are STICKY: the first non-zero status word poisons the handle (its state was only
// partially advanced), every later step / get / sync / reduce returns the same error; only nls_destroy is valid.
int poisoned(nls_engine *e) {
  switch (e->poison) {
  case 0: return NLS_OK;
  case 2: return fail(NLS_ERR_CUDA, "z-slab exchange timed out waiting for a peer rank (handle poisoned)");
  case 3: return fail(NLS_ERR_CUDA, "TMA pipeline timed out (internal error; handle poisoned)");
  case 4: return fail(NLS_ERR_CUDA, "persistent kernel: grid barrier timed out or was aborted; the state is inconsistent (handle poisoned)");
  default: return fail(NLS_ERR_NONFINITE, "tridiagonal eigensolver did not converge (non-finite state?; handle poisoned)");
  }
}

// device -> dense device buffer, no synchronisation
int copy_out_async(nls_engine *e, void *dst, const void *dev, size_t esz, int nb) {
  if (e->g.stride == e->g.n) {
    CU(cudaMemcpyAsync(dst, dev, (size_t)nb * e->g.n * esz, cudaMemcpyDeviceToDevice, e->stream));
  } else {
    for (int b = 0; b < nb; ++b)
      CU(cudaMemcpyAsync((char *)dst + (size_t)b * e->g.n * esz, (const char *)dev + ((size_t)b * e->g.stride + e->g.halo) * esz,
                         e->g.n * esz, cudaMemcpyDeviceToDevice, e->stream));
  }
  return NLS_OK;
}

int check_status(nls_engine *e) {
  if (e->poison) return poisoned(e);
  int st = 0;
  CU(cudaMemcpyAsync(&st, e->status, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  e->poison = st;
  return poisoned(e);
}

} // namespace

// =========================================================================== C ABI
extern "C" {

const char *nls_last_error(void) { return g_err.c_str(); }
int nls_version(void) { return 100; }

int nls_device_count(void) {
  int n = 0;
  CU(cudaGetDeviceCount(&n));
  return n;
}

int nls_destroy(nls_handle e) {
  if (!e) return NLS_OK;
  cudaSetDevice(e->d.device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  if (e->d2h_stream) cudaStreamSynchronize(e->d2h_stream);
  if (e->h2d_stream) cudaStreamSynchronize(e->h2d_stream);
  for (auto &ge : e->gexec)
    if (ge) cudaGraphExecDestroy(ge);
  if (e->arena) { // sub-allocations of the exchange arena
    e->wbuf[0] = e->wbuf[1] = e->w0 = e->u[0] = e->u[1] = e->u[2] = e->V = nullptr;
    e->cf = nullptr;
    for (int q = 0; q < e->nranks; ++q)
      if (e->peers[q].ipc && e->peers[q].base) cudaIpcCloseMemHandle(e->peers[q].base);
  }
  void *ptrs[] = {e->V, e->wbuf[0], e->wbuf[1], e->w0, e->u[0], e->u[1], e->u[2], e->acc, e->mf, e->cf,
                  e->scalA, e->scalB, e->coef, e->redout, e->rc.partials, e->rc.tickets, e->status,
                  e->pk_part, e->pk_bar, e->sv2_prog, e->dft_wx, e->dft_wy, e->dft_t1, e->arena, e->slab_dev, e->slab_scratch, e->rs};
  for (void *p : ptrs)
    if (p) cudaFree(p);
  for (int k = 0; k < 2; ++k) {
    if (e->ring[k]) cudaFree(e->ring[k]);
    if (e->stage[k]) cudaFree(e->stage[k]);
    for (cudaEvent_t ev : {e->ring_ready[k], e->ring_done[k], e->stage_ready[k], e->stage_free[k]})
      if (ev) cudaEventDestroy(ev);
  }
  if (e->d2h_stream) cudaStreamDestroy(e->d2h_stream);
  if (e->h2d_stream) cudaStreamDestroy(e->h2d_stream);
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
  return NLS_OK;
}

int nls_create(const nls_desc *desc, nls_handle *out) {
  if (!desc || !out || desc->struct_size != sizeof(nls_desc)) return fail(NLS_ERR_ARG, "bad descriptor");
  const nls_desc &d = *desc;
  const bool slab_req = d.nz_global && d.nz_global != d.nz; // a slab may hold just 2 planes of a taller grid
  if (d.nx < 3 || d.ny < 3 || d.nz == 0 || (d.nz == 2 && !slab_req) || d.krylov_m < 2 || d.krylov_m > (uint32_t)kMaxM ||
      d.batch < 1)
    return fail(NLS_ERR_ARG, "bad sizes: need nx,ny>=3, nz==1 or >=3, 2<=m<=64, batch>=1");
  if (d.equation < NLS_EQ_NLSE_CUBIC || d.equation > NLS_EQ_SGE_SINH) return fail(NLS_ERR_ARG, "bad equation");
  const bool cplx = d.equation <= NLS_EQ_NLSE_SAT;
  if (cplx && d.integrator != NLS_INT_SS2 && d.integrator != NLS_INT_SEWI)
    return fail(NLS_ERR_ARG, "NLSE needs NLS_INT_SS2 or NLS_INT_SEWI");
  if (!cplx && d.integrator != NLS_INT_GAUTSCHI && d.integrator != NLS_INT_SV)
    return fail(NLS_ERR_ARG, "real equations need NLS_INT_GAUTSCHI or NLS_INT_SV");
  if (!(d.dx > 0) || !(d.dy > 0)) return fail(NLS_ERR_ARG, "dx, dy must be positive");
  const uint32_t nzg = d.nz_global ? d.nz_global : d.nz;
  if (d.nz_global && (d.nz == 1 || d.z0 + d.nz > nzg)) return fail(NLS_ERR_ARG, "bad slab");
  if (d.nz_global && nzg != d.nz && (d.batch != 1 || d.nz < 2))
    return fail(NLS_ERR_ARG, "z-slab handles need batch == 1 and at least 2 planes per slab");
  if ((unsigned long long)d.nx * d.ny * nzg >= (1ull << 32)) return fail(NLS_ERR_ARG, "grid too large (2^32 sites)");
  int ndev = 0;
  CU(cudaGetDeviceCount(&ndev));
  if (d.device < 0 || d.device >= ndev) return fail(NLS_ERR_CUDA, "no such CUDA device (there is no CPU fallback)");
  CU(cudaSetDevice(d.device));

  if (d.reorth < NLS_REORTH_FULL || d.reorth > NLS_REORTH_BLOCK) return fail(NLS_ERR_ARG, "bad reorth mode");
  nls_engine *e = new nls_engine;
  e->d = d;
  // NLS_REORTH_FULL on the multi-kernel path = the Gram-delayed form unless the handle asks for a pass variant that only
  // exists in the two-pass form; NLS_REORTH=0 in the environment keeps the two-pass form (A/B knob)
  e->gdr_auto = !(d.flags & (NLS_FLAG_TMA | NLS_FLAG_MARCH));
  if (const char *ro = getenv("NLS_REORTH")) e->gdr_auto = e->gdr_auto && atoi(ro) != 0;
  e->cplx = cplx;
  e->W = cplx ? 2 : 1;
  e->esz = cplx ? 16 : 8;
  e->m = (int)d.krylov_m;
  e->batch = (int)d.batch;
  Geom &g = e->g;
  g.nx = d.nx; g.ny = d.ny; g.nz_loc = d.nz; g.nz_glob = nzg; g.z0 = d.z0;
  g.dim3 = d.nz > 1;
  g.plane = g.dim3 ? (long long)d.nx * d.ny : 0;
  g.n = (long long)d.nx * d.ny * d.nz;
  g.nglob = (long long)d.nx * d.ny * nzg;
  g.goff = (long long)d.z0 * d.nx * d.ny;
  g.halo = (d.nz_global && nzg != d.nz) ? g.plane : 0;
  g.stride = g.n + 2 * g.halo;
  const double sgn = d.op_sign == 0.0 ? 1.0 : d.op_sign;
  // L *= 1/(dx*dy) (laplacians.hpp:49); 3D: 1/(dx*dx) (:153)
  g.sc = sgn * (1.0 / (g.dim3 ? d.dx * d.dx : d.dx * d.dy));
  // L2 prefetch lead of the flat passes in grid sweeps; NLS_PREFETCH / NLS_PREFETCH_FIRST are tuning knobs.  Measured
  // (profiles/r02l_bench_ab.jsonl): lead 1 takes C4 7.32 -> 6.35 ms/step, flat C3 10.62 -> 9.99, C5 (tensor-map pass A,
  // flat pass B) 50.5 -> 48.4; lead 2 is no better.  The stencil-carrying flat pass of a 3D grid is the exception
  // (flat C5 66.4 ms/step without, 72.1 with; 2D C4: 6.98 without, 6.33 with -- profiles/r02m_bench_ab.jsonl).
  g.pf = kDefaultPrefetch;
  g.pf_first = g.dim3 ? 0 : kDefaultPrefetch;
  // the recombination kernel is the exception too: its prefetch re-fetches a quarter of the columns from DRAM (C3: 4.60
  // GB read per launch with, 3.57 = algorithmic without; 855 -> 662 us; C3 9.60 -> 9.35, C5 48.4 -> 47.7 ms/step,
  // profiles/r02q_*): off by default, NLS_PREFETCH_FINISH=1 turns it on
  g.pf_fin = 0;
  if (const char *pf = getenv("NLS_PREFETCH")) g.pf = g.pf_first = atoi(pf);
  if (const char *pf = getenv("NLS_PREFETCH_FIRST")) g.pf_first = atoi(pf);
  if (const char *pf = getenv("NLS_PREFETCH_FINISH")) g.pf_fin = atoi(pf);
  if (g.pf_fin < 0 || g.pf_fin > 8) g.pf_fin = 0;
  if (g.pf < 0 || g.pf > 8) g.pf = 0;
  if (g.pf_first < 0 || g.pf_first > 8) g.pf_first = 0;
  cudaDeviceGetAttribute(&e->nsm, cudaDevAttrMultiProcessorCount, d.device);
  cudaDeviceGetAttribute(&e->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, d.device);
  e->smem_optin -= 4096; // dynamic budget: leave room for the kernels' static shared memory
  if (e->nsm < 1) e->nsm = 148;
  if (const char *fc = getenv("NLS_FIRST_CAP")) e->first_cap = atoi(fc); // tuning knob (2..12), not API
  long long want = (g.n + kThreads - 1) / kThreads;
  long long cap = 8LL * e->nsm; // at most 8 CTAs of 256 threads per SM
  e->nblk = (int)(want < cap ? want : cap);
  e->slab = g.halo != 0;
  e->slot_bytes = slot_bytes_of(g.stride, e->esz);
  e->colstride = e->slab ? e->slot_bytes / e->esz : e->vec_elems() + col_pad_bytes() / e->esz;
  e->nphase = cplx ? (d.integrator == NLS_INT_SEWI ? 3 : 2) : (d.integrator == NLS_INT_SV ? 3 : 1);

  auto alloc = [&](void **p, size_t bytes) -> bool {
    if (cudaMalloc(p, bytes) != cudaSuccess) return false;
    return cudaMemsetAsync(*p, 0, bytes, 0) == cudaSuccess;
  };
  bool ok = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) == cudaSuccess;
  const size_t vb = e->vec_bytes();
  const bool krylov = d.integrator != NLS_INT_SV;
  // the Krylov buffers are also used by nls_matfunc / nls_lanczos / nls_stencil_apply
  if (!(g.halo != 0)) ok = ok && alloc(&e->V, e->colstride * e->esz * (e->m - 1));
  if (e->slab) { // stencil inputs (incl. the basis columns) + mailbox in one arena the other ranks can map
    ok = ok && alloc((void **)&e->arena, kMailboxBytes + (size_t)(kArenaSlots + e->m - 1) * e->slot_bytes);
    if (ok) {
      char *b = e->arena + kMailboxBytes;
      e->wbuf[0] = b; e->wbuf[1] = b + e->slot_bytes; e->w0 = b + 2 * e->slot_bytes;
      e->u[0] = b + 3 * e->slot_bytes; e->u[1] = b + 4 * e->slot_bytes; e->u[2] = b + 5 * e->slot_bytes;
      e->cf = (double *)(b + 6 * e->slot_bytes);
      e->V = b + (size_t)kArenaSlots * e->slot_bytes;
    }
    ok = ok && alloc((void **)&e->slab_dev, sizeof(SlabCtx)) && alloc((void **)&e->slab_scratch, 64);
    ok = ok && alloc((void **)&e->mf, g.stride * 8);
  } else {
    ok = ok && alloc(&e->wbuf[0], vb) && alloc(&e->wbuf[1], vb) && alloc(&e->w0, vb);
    ok = ok && alloc(&e->u[0], vb) && alloc(&e->u[1], vb);
    if (d.integrator == NLS_INT_SV || d.integrator == NLS_INT_SEWI) ok = ok && alloc(&e->u[2], vb);
    ok = ok && alloc((void **)&e->mf, g.stride * 8) && alloc((void **)&e->cf, g.stride * 8);
  }
  if (d.integrator == NLS_INT_GAUTSCHI || d.integrator == NLS_INT_SEWI) ok = ok && alloc(&e->acc, vb);
  ok = ok && alloc((void **)&e->scalA, (size_t)e->batch * kScalStride * 8);
  ok = ok && alloc((void **)&e->scalB, (size_t)e->batch * kScalStride * 8);
  ok = ok && alloc((void **)&e->coef, (size_t)e->batch * 2 * kCoefStride * 8);
  ok = ok && alloc((void **)&e->redout, (size_t)e->batch * 8);
  ok = ok && alloc((void **)&e->rc.partials, (size_t)e->batch * e->nblk * 48 * 8);
  ok = ok && alloc((void **)&e->rc.tickets, (size_t)e->batch * 4);
  ok = ok && alloc((void **)&e->status, 4);
  (void)krylov;
  if (ok && d.laplacian == NLS_LAP_ANISO) { // default c == 1
    std::vector<double> ones(g.stride, 1.0);
    ok = cudaMemcpy(e->cf, ones.data(), g.stride * 8, cudaMemcpyHostToDevice) == cudaSuccess;
  }
  ok = ok && cudaDeviceSynchronize() == cudaSuccess;
  if (ok) pk_setup(e);
  if (ok) sv2_setup(e);
  if (ok) tma_setup(e);
  if (ok) tm_setup(e);
  if (!ok) {
    std::string msg = std::string("allocation failed: ") + cudaGetErrorString(cudaGetLastError());
    nls_destroy(e);
    return fail(NLS_ERR_CUDA, msg);
  }
  *out = e;
  return NLS_OK;
}

static int set_field_impl(nls_handle e, int field, const void *src, size_t bytes, bool dev, bool sync = true) {
  if (!e || !src) return fail(NLS_ERR_ARG, "null argument");
  if (e->poison) return poisoned(e);
  CU(cudaSetDevice(e->d.device));
  const size_t sb = (size_t)e->batch * e->g.n * e->esz, fb = (size_t)e->g.n * 8;
  int rc;
  switch (field) {
  case NLS_FIELD_U:
    if (bytes != sb) return fail(NLS_ERR_ARG, "set U: size mismatch");
    rc = copy_in(e, e->cplx ? e->u[e->phase] : (void *)cur_u(e), src, e->esz, e->batch, dev);
    break;
  case NLS_FIELD_U_PAST:
    if (bytes != sb) return fail(NLS_ERR_ARG, "set U_PAST: size mismatch");
    if (e->cplx) {
      if (e->d.integrator != NLS_INT_SEWI) return fail(NLS_ERR_STATE, "U_PAST: real equations or NLS_INT_SEWI only");
      rc = copy_in(e, e->u[(e->phase + 2) % 3], src, 16, e->batch, dev);
      if (rc) return rc;
      if (sync) CU(cudaStreamSynchronize(e->stream));
      e->sewi_started = true; // explicit u-: no SS2 start-up step
      return NLS_OK;
    }
    rc = copy_in(e, cur_upast(e), src, 8, e->batch, dev);
    break;
  case NLS_FIELD_V: { // u_past = u0 - dt*v0  (sg_single_driver.cpp:70)
    if (e->cplx) return fail(NLS_ERR_STATE, "V: real equations only");
    if (bytes != sb) return fail(NLS_ERR_ARG, "set V: size mismatch");
    rc = copy_in(e, e->wbuf[0], src, 8, e->batch, dev);
    if (rc) return rc;
    LAUNCH(e, k_upast_from_v, e->grid(), e->g, cur_u(e), (const double *)e->wbuf[0], cur_upast(e), e->d.dt);
    CU(cudaGetLastError());
    break;
  }
  case NLS_FIELD_M:
    if (bytes != fb) return fail(NLS_ERR_ARG, "set M: size mismatch");
    rc = copy_in(e, e->mf, src, 8, 1, dev);
    break;
  case NLS_FIELD_C:
    if (bytes != fb) return fail(NLS_ERR_ARG, "set C: size mismatch");
    rc = copy_in(e, e->cf, src, 8, 1, dev);
    break;
  default: return fail(NLS_ERR_ARG, "bad field");
  }
  if (rc) return rc;
  if (field == NLS_FIELD_U) e->sewi_started = false;
  if (e->slab && e->connected) { // collective on slab handles: refresh the neighbours' ghost planes
    if (field == NLS_FIELD_C) halo_exchange<double>(e, e->cf);
    else if (field == NLS_FIELD_U && !e->cplx) halo_exchange<double>(e, cur_u(e));
    CU(cudaGetLastError());
  }
  e->start_valid = false;
  if (sync) CU(cudaStreamSynchronize(e->stream)); // host buffer may be reused by the caller
  return NLS_OK;
}

static int get_field_impl(nls_handle e, int field, void *dst, size_t bytes, bool dev) {
  if (!e || !dst) return fail(NLS_ERR_ARG, "null argument");
  CU(cudaSetDevice(e->d.device));
  const size_t sb = (size_t)e->batch * e->g.n * e->esz;
  if (bytes != sb) return fail(NLS_ERR_ARG, "get: size mismatch");
  int rc;
  switch (field) {
  case NLS_FIELD_U:
    rc = copy_out(e, dst, e->cplx ? e->u[e->phase] : (void *)cur_u(e), e->esz, e->batch, dev);
    break;
  case NLS_FIELD_U_PAST:
    if (e->cplx) {
      if (e->d.integrator != NLS_INT_SEWI) return fail(NLS_ERR_STATE, "U_PAST: real equations or NLS_INT_SEWI only");
      rc = copy_out(e, dst, e->u[(e->phase + 2) % 3], 16, e->batch, dev);
      break;
    }
    rc = copy_out(e, dst, cur_upast(e), 8, e->batch, dev);
    break;
  case NLS_FIELD_V: // v = (u - u_past)/dt  (sg_single_driver.cpp:113)
    if (e->cplx) return fail(NLS_ERR_STATE, "V: real equations only");
    LAUNCH(e, k_velocity, e->grid(), e->g, cur_u(e), cur_upast(e), (double *)e->wbuf[0], e->d.dt);
    CU(cudaGetLastError());
    rc = copy_out(e, dst, e->wbuf[0], 8, e->batch, dev);
    break;
  default: return fail(NLS_ERR_ARG, "bad field");
  }
  if (rc) return rc;
  return check_status(e);
}

static int resample_check(nls_handle e, int field, uint32_t tnx, uint32_t tny, uint32_t tnz, size_t bytes) {
  if (e->slab) return fail(NLS_ERR_STATE, "nls_get_field_resampled is not available on z-slab handles");
  if (field != NLS_FIELD_U && field != NLS_FIELD_V) return fail(NLS_ERR_ARG, "resampled: NLS_FIELD_U or NLS_FIELD_V");
  if (field == NLS_FIELD_V && e->cplx) return fail(NLS_ERR_STATE, "V: real equations only");
  if (field == NLS_FIELD_V && !(e->d.dt > 0)) return fail(NLS_ERR_STATE, "V needs desc.dt > 0");
  const uint32_t nz = e->g.dim3 ? (uint32_t)e->g.nz_loc : 1u;
  if (tnx < 2 || tny < 2 || (nz == 1 ? tnz != 1 : tnz < 2) || tnx > (1u << 20) || tny > (1u << 20) || tnz > (1u << 20))
    return fail(NLS_ERR_ARG, "resampled: target extents must be >= 2 (tnz == 1 for 2D handles)");
  const size_t need = (size_t)tnx * tny * tnz * e->batch * e->esz;
  if (bytes != need) return fail(NLS_ERR_ARG, "resampled: size mismatch");
  return NLS_OK;
}
static int resample_launch(nls_handle e, int field, uint32_t tnx, uint32_t tny, uint32_t tnz, void *out) {
  const size_t tot = (size_t)tnx * tny * tnz;
  const unsigned gx = (unsigned)std::min<size_t>((tot + kThreads - 1) / kThreads, (size_t)e->nsm * 8);
  const dim3 gr(gx, (unsigned)e->batch);
  if (e->cplx)
    k_resample<double2, false><<<gr, kThreads, 0, e->stream>>>(e->g, (const double2 *)e->u[e->phase], nullptr, 0.0,
                                                               (int)tnx, (int)tny, (int)tnz, (double2 *)out);
  else if (field == NLS_FIELD_U)
    k_resample<double, false><<<gr, kThreads, 0, e->stream>>>(e->g, (const double *)cur_u(e), nullptr, 0.0, (int)tnx,
                                                              (int)tny, (int)tnz, (double *)out);
  else
    k_resample<double, true><<<gr, kThreads, 0, e->stream>>>(e->g, (const double *)cur_u(e), (const double *)cur_upast(e),
                                                             1.0 / e->d.dt, (int)tnx, (int)tny, (int)tnz, (double *)out);
  e->count();
  CU(cudaGetLastError());
  return NLS_OK;
}

int nls_get_field_resampled(nls_handle e, int field, uint32_t tnx, uint32_t tny, uint32_t tnz, void *dst, size_t bytes) {
  if (!e || !dst) return fail(NLS_ERR_ARG, "null argument");
  int rc = resample_check(e, field, tnx, tny, tnz, bytes);
  if (rc) return rc;
  CU(cudaSetDevice(e->d.device));
  if (e->rs_bytes < bytes) {
    if (e->rs) CU(cudaFree(e->rs));
    e->rs = nullptr; e->rs_bytes = 0;
    CU(cudaMalloc(&e->rs, bytes));
    e->rs_bytes = bytes;
  }
  rc = resample_launch(e, field, tnx, tny, tnz, e->rs);
  if (rc) return rc;
  CU(cudaMemcpyAsync(dst, e->rs, bytes, cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return check_status(e);
}

// ---- spectral downsampling: downsample_fft (finalized_scripts/downsampling.py:10-34) as a direct cropped DFT
static int dft_check(nls_handle e, int field, uint32_t tnx, uint32_t tny, size_t bytes) {
  if (e->slab || e->g.dim3) return fail(NLS_ERR_STATE, "fft_resampled: 2D single-GPU handles only (the script's downsample_fft is 2D)");
  if (field != NLS_FIELD_U && field != NLS_FIELD_V) return fail(NLS_ERR_ARG, "fft_resampled: NLS_FIELD_U or NLS_FIELD_V");
  if (field == NLS_FIELD_V && e->cplx) return fail(NLS_ERR_STATE, "V: real equations only");
  if (field == NLS_FIELD_V && !(e->d.dt > 0)) return fail(NLS_ERR_STATE, "V needs desc.dt > 0");
  if (tnx < 1 || tny < 1 || tnx > (uint32_t)e->g.nx || tny > (uint32_t)e->g.ny)
    return fail(NLS_ERR_ARG, "fft_resampled: need 1 <= tnx <= nx and 1 <= tny <= ny");
  if (bytes != (size_t)tnx * tny * e->batch * e->esz) return fail(NLS_ERR_ARG, "fft_resampled: size mismatch");
  return NLS_OK;
}
static int dft_launch(nls_handle e, int field, uint32_t tnx, uint32_t tny, void *out) {
  if (e->dft_tx != tnx || e->dft_ty != tny) { // (re)build the two weight matrices for this target shape
    CU(cudaStreamSynchronize(e->stream));
    cudaFree(e->dft_wx); cudaFree(e->dft_wy); cudaFree(e->dft_t1);
    e->dft_wx = e->dft_wy = e->dft_t1 = nullptr;
    e->dft_tx = e->dft_ty = 0;
    CU(cudaMalloc((void **)&e->dft_wx, (size_t)tnx * e->g.nx * 16));
    CU(cudaMalloc((void **)&e->dft_wy, (size_t)tny * e->g.ny * 16));
    CU(cudaMalloc((void **)&e->dft_t1, (size_t)e->batch * tny * e->g.nx * 16));
    k_dft_weights<<<e->nsm * 4, kThreads, 0, e->stream>>>(e->dft_wx, e->g.nx, (int)tnx);
    k_dft_weights<<<e->nsm * 4, kThreads, 0, e->stream>>>(e->dft_wy, e->g.ny, (int)tny);
    e->launches += 2;
    CU(cudaGetLastError());
    e->dft_tx = tnx; e->dft_ty = tny;
  }
  const dim3 g1((e->g.nx + kThreads - 1) / kThreads, (unsigned)e->batch, (tny + 3) / 4);
  const dim3 g2((unsigned)std::min<size_t>(((size_t)tnx * tny + kThreads - 1) / kThreads, (size_t)e->nsm * 8), (unsigned)e->batch);
  if (e->cplx) {
    k_dft_cols<double2, false><<<g1, kThreads, 0, e->stream>>>(e->g, (const double2 *)e->u[e->phase], nullptr, 0.0, e->dft_wy, (int)tny, e->dft_t1);
    k_dft_rows<double2><<<g2, kThreads, 0, e->stream>>>(e->dft_t1, e->dft_wx, e->g.nx, (int)tnx, (int)tny, (double2 *)out);
  } else {
    if (field == NLS_FIELD_U)
      k_dft_cols<double, false><<<g1, kThreads, 0, e->stream>>>(e->g, (const double *)cur_u(e), nullptr, 0.0, e->dft_wy, (int)tny, e->dft_t1);
    else
      k_dft_cols<double, true><<<g1, kThreads, 0, e->stream>>>(e->g, (const double *)cur_u(e), (const double *)cur_upast(e), 1.0 / e->d.dt,
                                                               e->dft_wy, (int)tny, e->dft_t1);
    k_dft_rows<double><<<g2, kThreads, 0, e->stream>>>(e->dft_t1, e->dft_wx, e->g.nx, (int)tnx, (int)tny, (double *)out);
  }
  e->launches += 2;
  CU(cudaGetLastError());
  return NLS_OK;
}

int nls_get_field_fft_resampled(nls_handle e, int field, uint32_t tnx, uint32_t tny, void *dst, size_t bytes) {
  if (!e || !dst) return fail(NLS_ERR_ARG, "null argument");
  if (e->poison) return poisoned(e);
  int rc = dft_check(e, field, tnx, tny, bytes);
  if (rc) return rc;
  CU(cudaSetDevice(e->d.device));
  if (e->rs_bytes < bytes) {
    if (e->rs) CU(cudaFree(e->rs));
    e->rs = nullptr; e->rs_bytes = 0;
    CU(cudaMalloc(&e->rs, bytes));
    e->rs_bytes = bytes;
  }
  rc = dft_launch(e, field, tnx, tny, e->rs);
  if (rc) return rc;
  CU(cudaMemcpyAsync(dst, e->rs, bytes, cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return check_status(e);
}

// ---- on-device field / IC generators
int nls_generate_field(nls_handle e, int field, int32_t traj, const nls_gen_term *terms, uint32_t nterms, int post,
                       const double *pp, double Lx, double Ly, double Lz) {
  if (!e) return fail(NLS_ERR_ARG, "null handle");
  if (e->poison) return poisoned(e);
  if (post < NLS_GENPOST_NONE || post > NLS_GENPOST_GRADIENT_MASS) return fail(NLS_ERR_ARG, "generate: bad post-operation");
  if (post != NLS_GENPOST_GRADIENT_MASS && (!terms || nterms == 0 || nterms > 4096)) return fail(NLS_ERR_ARG, "generate: need 1..4096 terms");
  if (post != NLS_GENPOST_NONE && !pp) return fail(NLS_ERR_ARG, "generate: post-operation parameters missing");
  if (field != NLS_FIELD_U && field != NLS_FIELD_M && field != NLS_FIELD_C) return fail(NLS_ERR_ARG, "generate: NLS_FIELD_U, _M or _C");
  if (field == NLS_FIELD_U && (traj < -1 || traj >= e->batch)) return fail(NLS_ERR_ARG, "generate: bad trajectory index");
  if ((post == NLS_GENPOST_MINMAX || post == NLS_GENPOST_MAXABS) && e->slab)
    return fail(NLS_ERR_STATE, "generate: normalising post-operations are not available on z-slab handles");
  if (post == NLS_GENPOST_GRADIENT_MASS && (e->d.laplacian != NLS_LAP_ANISO || e->g.dim3 || field != NLS_FIELD_M))
    return fail(NLS_ERR_STATE, "generate: GRADIENT_MASS fills m(x) from the c field of a 2D div(c grad) handle");
  for (uint32_t q = 0; q < nterms; ++q)
    if (terms[q].kind < NLS_GEN_CONST || terms[q].kind > NLS_GEN_KINK) return fail(NLS_ERR_ARG, "generate: bad term kind");
  CU(cudaSetDevice(e->d.device));
  nls_gen_term *dterms = nullptr;
  unsigned long long *mm = nullptr;
  auto cleanup = [&]() { cudaFree(dterms); cudaFree(mm); };
  const int nb = (field == NLS_FIELD_U && traj < 0) ? e->batch : 1;
  if (nterms) {
    if (cudaMalloc(&dterms, nterms * sizeof(nls_gen_term)) != cudaSuccess) return fail(NLS_ERR_CUDA, "generate: allocation failed");
    cudaMemcpyAsync(dterms, terms, nterms * sizeof(nls_gen_term), cudaMemcpyHostToDevice, e->stream);
  }
  std::vector<unsigned long long> init(3 * nb);
  for (int b = 0; b < nb; ++b) { init[3 * b] = ~0ull; init[3 * b + 1] = 0ull; init[3 * b + 2] = 0ull; }
  if (cudaMalloc(&mm, init.size() * 8) != cudaSuccess) { cleanup(); return fail(NLS_ERR_CUDA, "generate: allocation failed"); }
  cudaMemcpyAsync(mm, init.data(), init.size() * 8, cudaMemcpyHostToDevice, e->stream);
  const double p0 = pp ? pp[0] : 0.0, p1 = pp ? pp[1] : 0.0, p2 = pp ? pp[2] : 0.0;
  const dim3 gr((unsigned)e->nblk, (unsigned)nb);
  const bool cstate = field == NLS_FIELD_U && e->cplx;
  void *dst;
  if (field == NLS_FIELD_M) dst = e->mf;
  else if (field == NLS_FIELD_C) dst = e->cf;
  else dst = (char *)(e->cplx ? e->u[e->phase] : (void *)cur_u(e)) + (traj > 0 ? (size_t)traj * e->g.stride * e->esz : 0);
  if (cstate) {
    k_generate<double2><<<gr, kThreads, 0, e->stream>>>(e->g, dterms, (int)nterms, post, p0, p1, p2, e->cf, Lx, Ly, Lz, (double2 *)dst,
                                                     e->g.stride, mm);
    if (post == NLS_GENPOST_MINMAX || post == NLS_GENPOST_MAXABS)
      k_generate_post<double2><<<gr, kThreads, 0, e->stream>>>(e->g, post, p0, (double2 *)dst, e->g.stride, mm);
  } else {
    k_generate<double><<<gr, kThreads, 0, e->stream>>>(e->g, dterms, (int)nterms, post, p0, p1, p2, e->cf, Lx, Ly, Lz, (double *)dst,
                                                    e->g.stride, mm);
    if (post == NLS_GENPOST_MINMAX || post == NLS_GENPOST_MAXABS)
      k_generate_post<double><<<gr, kThreads, 0, e->stream>>>(e->g, post, p0, (double *)dst, e->g.stride, mm);
  }
  e->launches += (post == NLS_GENPOST_MINMAX || post == NLS_GENPOST_MAXABS) ? 2 : 1;
  cudaError_t ce = cudaGetLastError();
  if (ce == cudaSuccess && field == NLS_FIELD_U) e->sewi_started = false;
  if (ce == cudaSuccess && e->slab && e->connected) { // collective on slab handles, like nls_set_field
    if (field == NLS_FIELD_C) halo_exchange<double>(e, e->cf);
    else if (field == NLS_FIELD_U && !e->cplx) halo_exchange<double>(e, cur_u(e));
    ce = cudaGetLastError();
  }
  e->start_valid = false;
  if (ce == cudaSuccess) ce = cudaStreamSynchronize(e->stream); // terms / mm are freed below
  cleanup();
  if (ce != cudaSuccess) return fail(NLS_ERR_CUDA, std::string("generate: ") + cudaGetErrorString(ce));
  return NLS_OK;
}

// ---- asynchronous transfers (replaces the blocking cudaMemcpy per snapshot of nlse_dev.hpp:323-334)
static int async_init(nls_handle e) {
  if (e->async_used) return NLS_OK;
  CU(cudaStreamCreateWithFlags(&e->d2h_stream, cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&e->h2d_stream, cudaStreamNonBlocking));
  for (int k = 0; k < 2; ++k) {
    CU(cudaEventCreateWithFlags(&e->ring_ready[k], cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&e->ring_done[k], cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&e->stage_ready[k], cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&e->stage_free[k], cudaEventDisableTiming));
  }
  e->async_used = true;
  return NLS_OK;
}
static int ring_slot(nls_handle e, size_t bytes, int *slot) {
  const int k = e->ring_next;
  e->ring_next ^= 1;
  CU(cudaStreamWaitEvent(e->stream, e->ring_done[k], 0)); // the previous D2H out of this slot has finished
  if (e->ring_bytes[k] < bytes) {
    CU(cudaStreamSynchronize(e->d2h_stream));
    if (e->ring[k]) CU(cudaFree(e->ring[k]));
    e->ring[k] = nullptr; e->ring_bytes[k] = 0;
    CU(cudaMalloc(&e->ring[k], bytes));
    e->ring_bytes[k] = bytes;
  }
  *slot = k;
  return NLS_OK;
}
// snapshot in ring[k] (dense [batch][per] bytes) -> host, on the D2H stream, overlapping whatever the compute stream does next
static int ring_drain(nls_handle e, int k, void *dst, size_t per, size_t dst_stride) {
  CU(cudaEventRecord(e->ring_ready[k], e->stream));
  CU(cudaStreamWaitEvent(e->d2h_stream, e->ring_ready[k], 0));
  if (dst_stride == 0 || dst_stride == per || e->batch == 1) {
    CU(cudaMemcpyAsync(dst, e->ring[k], per * e->batch, cudaMemcpyDeviceToHost, e->d2h_stream));
  } else {
    for (int b = 0; b < e->batch; ++b)
      CU(cudaMemcpyAsync((char *)dst + (size_t)b * dst_stride, (const char *)e->ring[k] + (size_t)b * per, per,
                         cudaMemcpyDeviceToHost, e->d2h_stream));
  }
  CU(cudaEventRecord(e->ring_done[k], e->d2h_stream));
  return NLS_OK;
}

int nls_get_field_async(nls_handle e, int field, void *dst, size_t bytes, size_t dst_traj_stride) {
  if (!e || !dst) return fail(NLS_ERR_ARG, "null argument");
  if (e->poison) return poisoned(e);
  CU(cudaSetDevice(e->d.device));
  const size_t per = (size_t)e->g.n * e->esz, sb = per * e->batch;
  if (bytes != sb) return fail(NLS_ERR_ARG, "get_async: size mismatch");
  if (dst_traj_stride != 0 && dst_traj_stride < per) return fail(NLS_ERR_ARG, "get_async: trajectory stride smaller than one state");
  const void *src;
  switch (field) {
  case NLS_FIELD_U: src = e->cplx ? e->u[e->phase] : (void *)cur_u(e); break;
  case NLS_FIELD_U_PAST:
    if (e->cplx && e->d.integrator != NLS_INT_SEWI) return fail(NLS_ERR_STATE, "U_PAST: real equations or NLS_INT_SEWI only");
    src = e->cplx ? e->u[(e->phase + 2) % 3] : (void *)cur_upast(e);
    break;
  case NLS_FIELD_V:
    if (e->cplx) return fail(NLS_ERR_STATE, "V: real equations only");
    src = nullptr;
    break;
  default: return fail(NLS_ERR_ARG, "bad field");
  }
  int rc = async_init(e);
  if (rc) return rc;
  int k;
  rc = ring_slot(e, sb, &k);
  if (rc) return rc;
  if (field == NLS_FIELD_V) { // v = (u - u_past)/dt straight into the ring (dense layout: halo-free view)
    Geom gd = e->g;
    if (e->g.halo == 0) {
      LAUNCH(e, k_velocity, e->grid(), gd, cur_u(e), cur_upast(e), (double *)e->ring[k], e->d.dt);
      CU(cudaGetLastError());
    } else {
      LAUNCH(e, k_velocity, e->grid(), e->g, cur_u(e), cur_upast(e), (double *)e->wbuf[0], e->d.dt);
      CU(cudaGetLastError());
      rc = copy_out_async(e, e->ring[k], e->wbuf[0], 8, e->batch);
      if (rc) return rc;
    }
  } else {
    rc = copy_out_async(e, e->ring[k], src, e->esz, e->batch);
    if (rc) return rc;
  }
  return ring_drain(e, k, dst, per, dst_traj_stride);
}

int nls_get_field_resampled_async(nls_handle e, int field, uint32_t tnx, uint32_t tny, uint32_t tnz, void *dst, size_t bytes,
                                  size_t dst_traj_stride) {
  if (!e || !dst) return fail(NLS_ERR_ARG, "null argument");
  if (e->poison) return poisoned(e);
  int rc = resample_check(e, field, tnx, tny, tnz, bytes);
  if (rc) return rc;
  CU(cudaSetDevice(e->d.device));
  const size_t per = (size_t)tnx * tny * tnz * e->esz;
  if (dst_traj_stride != 0 && dst_traj_stride < per) return fail(NLS_ERR_ARG, "resampled_async: trajectory stride too small");
  rc = async_init(e);
  if (rc) return rc;
  int k;
  rc = ring_slot(e, bytes, &k);
  if (rc) return rc;
  rc = resample_launch(e, field, tnx, tny, tnz, e->ring[k]);
  if (rc) return rc;
  return ring_drain(e, k, dst, per, dst_traj_stride);
}

int nls_get_field_fft_resampled_async(nls_handle e, int field, uint32_t tnx, uint32_t tny, void *dst, size_t bytes,
                                      size_t dst_traj_stride) {
  if (!e || !dst) return fail(NLS_ERR_ARG, "null argument");
  if (e->poison) return poisoned(e);
  int rc = dft_check(e, field, tnx, tny, bytes);
  if (rc) return rc;
  CU(cudaSetDevice(e->d.device));
  const size_t per = (size_t)tnx * tny * e->esz;
  if (dst_traj_stride != 0 && dst_traj_stride < per) return fail(NLS_ERR_ARG, "fft_resampled_async: trajectory stride too small");
  rc = async_init(e);
  if (rc) return rc;
  int k;
  rc = ring_slot(e, bytes, &k);
  if (rc) return rc;
  rc = dft_launch(e, field, tnx, tny, e->ring[k]);
  if (rc) return rc;
  return ring_drain(e, k, dst, per, dst_traj_stride);
}

// host (ideally page-locked) -> staging buffer on the H2D stream, then a device-to-device copy on the compute
// stream: the upload overlaps the steps still running, the caller's buffer is free again after nls_wait
int nls_set_field_async(nls_handle e, int field, const void *src, size_t bytes) {
  if (!e || !src) return fail(NLS_ERR_ARG, "null argument");
  if (e->poison) return poisoned(e);
  if (field == NLS_FIELD_V) return fail(NLS_ERR_ARG, "set_async: U, U_PAST, M or C");
  CU(cudaSetDevice(e->d.device));
  int rc = async_init(e);
  if (rc) return rc;
  const int k = e->stage_next;
  e->stage_next ^= 1;
  CU(cudaStreamWaitEvent(e->h2d_stream, e->stage_free[k], 0)); // the compute stream has consumed this slot
  if (e->stage_bytes[k] < bytes) {
    CU(cudaStreamSynchronize(e->stream));
    CU(cudaStreamSynchronize(e->h2d_stream));
    if (e->stage[k]) CU(cudaFree(e->stage[k]));
    e->stage[k] = nullptr; e->stage_bytes[k] = 0;
    CU(cudaMalloc(&e->stage[k], bytes));
    e->stage_bytes[k] = bytes;
  }
  CU(cudaMemcpyAsync(e->stage[k], src, bytes, cudaMemcpyHostToDevice, e->h2d_stream));
  CU(cudaEventRecord(e->stage_ready[k], e->h2d_stream));
  CU(cudaStreamWaitEvent(e->stream, e->stage_ready[k], 0));
  rc = set_field_impl(e, field, e->stage[k], bytes, true, false);
  if (rc) return rc;
  CU(cudaEventRecord(e->stage_free[k], e->stream));
  return NLS_OK;
}

// page-locked host memory for the asynchronous transfers (callers without a CUDA runtime of their own)
int nls_host_alloc(void **out, size_t bytes) {
  if (!out) return fail(NLS_ERR_ARG, "null argument");
  CU(cudaMallocHost(out, bytes));
  return NLS_OK;
}
int nls_host_free(void *p) {
  if (p) CU(cudaFreeHost(p));
  return NLS_OK;
}
int nls_host_register(void *p, size_t bytes) {
  if (!p) return fail(NLS_ERR_ARG, "null argument");
  CU(cudaHostRegister(p, bytes, cudaHostRegisterDefault));
  return NLS_OK;
}
int nls_host_unregister(void *p) {
  if (p) CU(cudaHostUnregister(p));
  return NLS_OK;
}

// A page-locked source (cudaMallocHost / cudaHostRegister / nls_host_alloc) is copied asynchronously: the call
// returns at once and the buffer must stay unchanged until the next nls_sync / nls_wait / blocking get on this handle.
int nls_set_field(nls_handle h, int f, const void *s, size_t b) {
  bool pinned = false;
  if (s) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, s) == cudaSuccess) pinned = at.type == cudaMemoryTypeHost;
    else cudaGetLastError();
  }
  return set_field_impl(h, f, s, b, false, !pinned);
}
int nls_set_field_dev(nls_handle h, int f, const void *s, size_t b) { return set_field_impl(h, f, s, b, true, false); }
int nls_get_field(nls_handle h, int f, void *d, size_t b) { return get_field_impl(h, f, d, b, false); }
int nls_get_field_dev(nls_handle h, int f, void *d, size_t b) { return get_field_impl(h, f, d, b, true); }

static int slab_ready(nls_handle e) {
  if (e->slab && !e->connected) return fail(NLS_ERR_STATE, "z-slab handle: call nls_slab_connect first");
  return NLS_OK;
}

int nls_step(nls_handle e, double dt, uint32_t nsteps) {
  if (!e) return fail(NLS_ERR_ARG, "null handle");
  if (slab_ready(e)) return NLS_ERR_STATE;
  if (e->poison) return poisoned(e);
  CU(cudaSetDevice(e->d.device));
  return run_steps(e, dt, nsteps);
}

int nls_sync(nls_handle e) {
  if (!e) return fail(NLS_ERR_ARG, "null handle");
  CU(cudaSetDevice(e->d.device));
  CU(cudaStreamSynchronize(e->stream));
  if (e->async_used) { // outstanding nls_get_field_async / nls_set_field_async transfers
    CU(cudaStreamSynchronize(e->h2d_stream));
    CU(cudaStreamSynchronize(e->d2h_stream));
  }
  return check_status(e);
}

int nls_wait(nls_handle e) { return nls_sync(e); }

int nls_step_timed(nls_handle e, double dt, uint32_t nsteps, float *ms) {
  if (!e || !ms) return fail(NLS_ERR_ARG, "null argument");
  if (slab_ready(e)) return NLS_ERR_STATE;
  CU(cudaSetDevice(e->d.device));
  int rc = ensure_start(e, dt);
  if (rc) return rc;
  cudaEvent_t a, b;
  CU(cudaEventCreate(&a));
  CU(cudaEventCreate(&b));
  CU(cudaEventRecord(a, e->stream));
  rc = run_steps(e, dt, nsteps);
  CU(cudaEventRecord(b, e->stream));
  CU(cudaEventSynchronize(b));
  CU(cudaEventElapsedTime(ms, a, b));
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  return rc ? rc : check_status(e);
}

uint64_t nls_launch_count(nls_handle e) { return e ? e->launches : 0; }
uint64_t nls_tma_launch_count(nls_handle e) { return e ? e->tma_launches : 0; }
uint64_t nls_tmap_launch_count(nls_handle e) { return e ? e->tm_launches : 0; }
void *nls_stream(nls_handle e) { return e ? (void *)e->stream : nullptr; }

static int stencil_apply_impl(nls_handle e, const void *x, void *y, bool dev) {
  if (!e || !x || !y) return fail(NLS_ERR_ARG, "null argument");
  if (e->poison) return poisoned(e);
  CU(cudaSetDevice(e->d.device));
  if (slab_ready(e)) return NLS_ERR_STATE;
  int rc = copy_in(e, e->w0, x, e->esz, e->batch, dev);
  if (rc) return rc;
  if (e->cplx) halo_exchange<double2>(e, e->w0);
  else halo_exchange<double>(e, e->w0);
  const bool an = e->d.laplacian == NLS_LAP_ANISO;
  if (e->cplx) {
    if (an) LAUNCH(e, (k_stencil<double2, true>), e->grid(), e->g, (const double2 *)e->w0, (double2 *)e->wbuf[0], e->cf);
    else LAUNCH(e, (k_stencil<double2, false>), e->grid(), e->g, (const double2 *)e->w0, (double2 *)e->wbuf[0], e->cf);
  } else {
    if (an) LAUNCH(e, (k_stencil<double, true>), e->grid(), e->g, (const double *)e->w0, (double *)e->wbuf[0], e->cf);
    else LAUNCH(e, (k_stencil<double, false>), e->grid(), e->g, (const double *)e->w0, (double *)e->wbuf[0], e->cf);
  }
  // slabs: trailing cross-slab barrier so a peer's next upload cannot overwrite ghost planes still being read
  if (e->cplx) halo_exchange<double2>(e, e->wbuf[0]);
  else halo_exchange<double>(e, e->wbuf[0]);
  CU(cudaGetLastError());
  e->start_valid = false; // w0 was overwritten
  if (dev) return copy_out_async(e, y, e->wbuf[0], e->esz, e->batch); // stream-ordered, like the reference's multiply()
  return copy_out(e, y, e->wbuf[0], e->esz, e->batch, false);
}
int nls_stencil_apply(nls_handle e, const void *x, void *y) { return stencil_apply_impl(e, x, y, false); }
int nls_stencil_apply_dev(nls_handle e, const void *x, void *y) { return stencil_apply_impl(e, x, y, true); }

// upload `in` into w0, ||in||^2 -> scalA.nrm2[0], run the Lanczos process
static int lanczos_from_host(nls_handle e, const void *in, bool dev = false) {
  if (slab_ready(e)) return NLS_ERR_STATE;
  int rc = copy_in(e, e->w0, in, e->esz, e->batch, dev);
  if (rc) return rc;
  e->start_valid = false;
  if (e->cplx) halo_exchange<double2>(e, e->w0);
  else halo_exchange<double>(e, e->w0);
  if (e->cplx) {
    LAUNCH(e, (k_norm2<double2>), e->grid(), e->g, (const double2 *)e->w0, e->scalA + kOffNrm2, kScalStride, e->rc);
    lanczos<double2>(e, (const double2 *)e->w0, e->scalA);
  } else {
    LAUNCH(e, (k_norm2<double>), e->grid(), e->g, (const double *)e->w0, e->scalA + kOffNrm2, kScalStride, e->rc);
    lanczos<double>(e, (const double *)e->w0, e->scalA);
  }
  CU(cudaGetLastError());
  return NLS_OK;
}

static int matfunc_impl(nls_handle e, int fn, double t_re, double t_im, const void *in, void *out, bool dev) {
  if (!e || !in || !out) return fail(NLS_ERR_ARG, "null argument");
  if (e->poison) return poisoned(e);
  if (fn < NLS_FN_EXP || fn > NLS_FN_MOD_COSINE) return fail(NLS_ERR_ARG, "bad function id");
  if (!e->cplx && fn <= NLS_FN_SINC && fn == NLS_FN_EXP && t_im != 0.0)
    return fail(NLS_ERR_STATE, "complex exp needs a complex-state handle");
  CU(cudaSetDevice(e->d.device));
  int rc = lanczos_from_host(e, in, dev);
  if (rc) return rc;
  tridiag(e, e->scalA, 1, fn, t_re, t_im, 0, 0, 0);
  void *dst = e->wbuf[1]; // scratch: neither state nor part of the basis
  if (e->cplx)
    LAUNCH(e, (k_combine<double2>), e->grid(), e->g, (const double2 *)e->w0, (const double2 *)e->V, e->colstride, e->coef, 0, e->m, (double2 *)dst);
  else
    LAUNCH(e, (k_combine<double>), e->grid(), e->g, (const double *)e->w0, (const double *)e->V, e->colstride, e->coef, 0, e->m, (double *)dst);
  CU(cudaGetLastError());
  if (dev) return copy_out_async(e, out, dst, e->esz, e->batch); // stream-ordered; errors surface at the next nls_sync
  rc = copy_out(e, out, dst, e->esz, e->batch, false);
  return rc ? rc : check_status(e);
}
int nls_matfunc(nls_handle e, int fn, double t_re, double t_im, const void *in, void *out) {
  return matfunc_impl(e, fn, t_re, t_im, in, out, false);
}
int nls_matfunc_dev(nls_handle e, int fn, double t_re, double t_im, const void *in, void *out) {
  return matfunc_impl(e, fn, t_re, t_im, in, out, true);
}

// 1 = device (or managed) pointer, 0 = host pointer
int nls_pointer_is_device(const void *p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return 0; }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}
int nls_memcpy_to_host(void *dst, const void *dev_src, size_t bytes) {
  if (!dst || !dev_src) return fail(NLS_ERR_ARG, "null argument");
  CU(cudaMemcpy(dst, dev_src, bytes, cudaMemcpyDeviceToHost));
  return NLS_OK;
}

// rel_err = max_i |(CSR x - L x)_i| / max_i |(L x)_i| for a deterministic pseudo-random x (see k_csr_residual)
int nls_csr_residual(nls_handle e, const int *row_ptr, const int *col_ind, const void *values, uint32_t n, uint64_t nnz,
                     int values_complex, int on_device, double *rel_err) {
  if (!e || !row_ptr || !col_ind || !values || !rel_err) return fail(NLS_ERR_ARG, "null argument");
  if (e->poison) return poisoned(e);
  if (e->slab || e->batch != 1) return fail(NLS_ERR_STATE, "csr check: single-trajectory, single-GPU handles only");
  if ((long long)n != e->g.n) return fail(NLS_ERR_ARG, "csr check: n does not match the handle's grid");
  if (values_complex && !e->cplx) return fail(NLS_ERR_ARG, "csr check: complex values need a complex-state handle");
  CU(cudaSetDevice(e->d.device));
  const size_t vb = (size_t)nnz * (values_complex ? 16 : 8);
  int *rp = nullptr, *ci = nullptr;
  double *val = nullptr;
  unsigned long long *out = nullptr;
  const int *rpd = row_ptr, *cid = col_ind;
  const double *vd = (const double *)values;
  auto cleanup = [&]() { cudaFree(rp); cudaFree(ci); cudaFree(val); cudaFree(out); };
  if (!on_device) {
    if (cudaMalloc(&rp, (size_t)(n + 1) * 4) != cudaSuccess || cudaMalloc(&ci, (size_t)nnz * 4) != cudaSuccess ||
        cudaMalloc(&val, vb) != cudaSuccess) { cleanup(); return fail(NLS_ERR_CUDA, "csr check: allocation failed"); }
    cudaMemcpyAsync(rp, row_ptr, (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, e->stream);
    cudaMemcpyAsync(ci, col_ind, (size_t)nnz * 4, cudaMemcpyHostToDevice, e->stream);
    cudaMemcpyAsync(val, values, vb, cudaMemcpyHostToDevice, e->stream);
    rpd = rp; cid = ci; vd = val;
  }
  if (cudaMalloc(&out, 16) != cudaSuccess) { cleanup(); return fail(NLS_ERR_CUDA, "csr check: allocation failed"); }
  cudaMemsetAsync(out, 0, 16, e->stream);
  const bool an = e->d.laplacian == NLS_LAP_ANISO;
  const dim3 gr = e->grid();
  if (e->cplx) {
    k_fill_hash<double2><<<gr, kThreads, 0, e->stream>>>(e->g, (double2 *)e->w0, 0x5eedull);
    if (an) k_stencil<double2, true><<<gr, kThreads, 0, e->stream>>>(e->g, (const double2 *)e->w0, (double2 *)e->wbuf[0], e->cf);
    else k_stencil<double2, false><<<gr, kThreads, 0, e->stream>>>(e->g, (const double2 *)e->w0, (double2 *)e->wbuf[0], e->cf);
    if (values_complex)
      k_csr_residual<double2, true><<<gr, kThreads, 0, e->stream>>>(e->g.n, rpd, cid, vd, (const double2 *)e->w0, (const double2 *)e->wbuf[0], out);
    else
      k_csr_residual<double2, false><<<gr, kThreads, 0, e->stream>>>(e->g.n, rpd, cid, vd, (const double2 *)e->w0, (const double2 *)e->wbuf[0], out);
  } else {
    k_fill_hash<double><<<gr, kThreads, 0, e->stream>>>(e->g, (double *)e->w0, 0x5eedull);
    if (an) k_stencil<double, true><<<gr, kThreads, 0, e->stream>>>(e->g, (const double *)e->w0, (double *)e->wbuf[0], e->cf);
    else k_stencil<double, false><<<gr, kThreads, 0, e->stream>>>(e->g, (const double *)e->w0, (double *)e->wbuf[0], e->cf);
    k_csr_residual<double, false><<<gr, kThreads, 0, e->stream>>>(e->g.n, rpd, cid, vd, (const double *)e->w0, (const double *)e->wbuf[0], out);
  }
  e->launches += 3;
  e->start_valid = false;
  unsigned long long h[2] = {0, 0};
  cudaError_t ce = cudaGetLastError();
  if (ce == cudaSuccess) ce = cudaMemcpyAsync(h, out, 16, cudaMemcpyDeviceToHost, e->stream);
  if (ce == cudaSuccess) ce = cudaStreamSynchronize(e->stream);
  cleanup();
  if (ce != cudaSuccess) return fail(NLS_ERR_CUDA, std::string("csr check: ") + cudaGetErrorString(ce));
  double dmax, ymax;
  std::memcpy(&dmax, &h[0], 8);
  std::memcpy(&ymax, &h[1], 8);
  *rel_err = ymax > 0 ? dmax / ymax : (dmax > 0 ? 1e300 : 0.0);
  return NLS_OK;
}

int nls_lanczos(nls_handle e, const void *in, void *hV, void *halpha, double *hbeta) {
  if (!e || !in) return fail(NLS_ERR_ARG, "null argument");
  CU(cudaSetDevice(e->d.device));
  int rc = lanczos_from_host(e, in);
  if (rc) return rc;
  CU(cudaStreamSynchronize(e->stream));
  const int m = e->m;
  std::vector<double> sc(kScalStride);
  CU(cudaMemcpy(sc.data(), e->scalA, kScalStride * 8, cudaMemcpyDeviceToHost));
  if (hbeta)
    for (int j = 0; j < m; ++j) hbeta[j] = std::sqrt(sc[kOffNrm2 + j]);
  if (halpha) {
    std::memset(halpha, 0, (size_t)m * e->esz); // T(m-1,m-1) stays 0
    std::memcpy(halpha, &sc[kOffAlpha], (size_t)(m - 1) * e->esz);
  }
  if (hV) { // columns are kept un-normalised on the device: V.col(i) = column_i / beta_i (presentation only)
    const size_t nb = (size_t)e->g.n * e->esz;
    for (int jc = 0; jc < m; ++jc) {
      const char *src = jc == 0 ? (const char *)e->w0 : (const char *)e->V + (size_t)(jc - 1) * e->colstride * e->esz;
      char *dstc = (char *)hV + (size_t)jc * nb;
      CU(cudaMemcpy(dstc, src + e->g.halo * e->esz, nb, cudaMemcpyDeviceToHost));
      const double inv = 1.0 / std::sqrt(sc[kOffNrm2 + jc]);
      double *p = (double *)dstc;
      for (size_t i = 0; i < (size_t)e->g.n * e->W; ++i) p[i] *= inv;
    }
  }
  return NLS_OK;
}

int nls_reduce(nls_handle e, int what, double *out) {
  if (!e || !out) return fail(NLS_ERR_ARG, "null argument");
  if (e->poison) return poisoned(e);
  CU(cudaSetDevice(e->d.device));
  if (what != NLS_RED_MASS && what != NLS_RED_NONFINITE && what != NLS_RED_ENERGY) return fail(NLS_ERR_ARG, "bad reduction");
  if (what == NLS_RED_ENERGY) {
    if (e->slab) return fail(NLS_ERR_STATE, "NLS_RED_ENERGY is not available on z-slab handles");
    // x, y, z spacings as the scripts use them: compare_utils_complex_2d.py:136-137 (dx for x, dy for y);
    // energy_sge.py:7-8 hands dx to array axis 0 (= y) and dy to axis 1 (= x) in 2D
    const bool swap = !e->cplx && !e->g.dim3;
    const double h0 = swap ? e->d.dy : e->d.dx, h1 = swap ? e->d.dx : e->d.dy, h2 = e->d.dz;
    if (e->cplx)
      LAUNCH(e, k_reduce_energy<double2>, e->grid(), e->g, (const double2 *)e->u[e->phase], (const double *)nullptr, 0.0,
             h0, h1, h2, e->redout, e->rc);
    else {
      if (!(e->d.dt > 0)) return fail(NLS_ERR_STATE, "NLS_RED_ENERGY on a real state needs desc.dt > 0 (v = (u-u_past)/dt)");
      LAUNCH(e, k_reduce_energy<double>, e->grid(), e->g, (const double *)cur_u(e), (const double *)cur_upast(e),
             1.0 / e->d.dt, h0, h1, h2, e->redout, e->rc);
    }
  } else if (e->cplx) {
    const double2 *u = (const double2 *)e->u[e->phase];
    if (what == NLS_RED_MASS) LAUNCH(e, (k_reduce_state<double2, NLS_RED_MASS>), e->grid(), e->g, u, e->redout, e->rc);
    else LAUNCH(e, (k_reduce_state<double2, NLS_RED_NONFINITE>), e->grid(), e->g, u, e->redout, e->rc);
  } else {
    const double *u = cur_u(e);
    if (what == NLS_RED_MASS) LAUNCH(e, (k_reduce_state<double, NLS_RED_MASS>), e->grid(), e->g, u, e->redout, e->rc);
    else LAUNCH(e, (k_reduce_state<double, NLS_RED_NONFINITE>), e->grid(), e->g, u, e->redout, e->rc);
  }
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(out, e->redout, (size_t)e->batch * 8, cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  if (what == NLS_RED_MASS || what == NLS_RED_ENERGY) {
    double cell = e->g.dim3 ? e->d.dx * e->d.dy * e->d.dz : e->d.dx * e->d.dy;
    if (what == NLS_RED_ENERGY && !e->cplx) cell *= 0.5; // energy_sge.py:14
    for (int b = 0; b < e->batch; ++b) out[b] *= cell;
  }
  return NLS_OK;
}

// ---------------------------------------------------------------- allocation-free mass reduction (nls.h)
// Scratch of the grid reduction in a file-scope device array instead of the handle's cudaMalloc'ed RedCtx buffers:
// [0, kNaCtas) doubles = per-CTA partial sums, then kNaBatch unsigned completion tickets.  In HOST code the name
// g_noalloc_scratch is only a stand-in for the symbol: its device address comes from cudaGetSymbolAddress, once per
// device with that device current (noalloc_scratch_ptr), and is what the kernel is given.
} // extern "C"
namespace {
constexpr unsigned kNaCtas = NLS_NOALLOC_MAX_CTAS, kNaBatch = NLS_NOALLOC_MAX_BATCH;
static __device__ __align__(256) unsigned char g_noalloc_scratch[NLS_NOALLOC_SCRATCH_BYTES];
constexpr int kNaMaxDev = 64;
unsigned char *g_noalloc_ptr[kNaMaxDev] = {};
int g_noalloc_nsm[kNaMaxDev] = {};
std::mutex g_noalloc_mu;

// variant of k_reduce_state<S, NLS_RED_MASS> (nls_kernels.cuh) without Geom / handle: trajectory t at u + t * stride
template <class S>
__global__ void __launch_bounds__(kThreads)
k_reduce_mass_noalloc(const S *__restrict__ u, long long n, long long stride, double cell, double *dst, RedCtx rc) {
  const S *ub = u + (size_t)blockIdx.y * stride;
  double acc[1] = {0.0};
  for (long long idx = (long long)blockIdx.x * kThreads + threadIdx.x; idx < n; idx += (long long)gridDim.x * kThreads)
    acc[0] += Sc<S>::abs2(ub[idx]);
  acc[0] *= cell; // nls_reduce scales the total on the host; here every thread's partial sum is scaled (differs in the last bits only)
  grid_reduce<1>(acc, rc, dst + blockIdx.y);
}
} // namespace
extern "C" {

int nls_reduce_mass_noalloc(int device, const void *dev_u, uint64_t n_sites, uint32_t batch, uint64_t traj_stride,
                            int is_complex, double cell, double *dev_out, void *stream) {
  if (!dev_u || !dev_out) return fail(NLS_ERR_ARG, "nls_reduce_mass_noalloc: null argument");
  if (device < 0 || device >= kNaMaxDev) return fail(NLS_ERR_ARG, "nls_reduce_mass_noalloc: bad device ordinal");
  if (n_sites == 0 || batch == 0) return fail(NLS_ERR_ARG, "nls_reduce_mass_noalloc: empty input");
  if (batch > kNaBatch)
    return fail(NLS_ERR_ARG, "nls_reduce_mass_noalloc: batch " + std::to_string(batch) + " exceeds NLS_NOALLOC_MAX_BATCH = " +
                                 std::to_string(kNaBatch) + " (static scratch, no allocation)");
  if (n_sites > NLS_NOALLOC_MAX_SITES / batch)
    return fail(NLS_ERR_ARG, "nls_reduce_mass_noalloc: n_sites * batch = " + std::to_string(n_sites) + " * " + std::to_string(batch) +
                                 " exceeds NLS_NOALLOC_MAX_SITES = " + std::to_string(NLS_NOALLOC_MAX_SITES));
  if (traj_stride == 0) traj_stride = n_sites;
  if (traj_stride < n_sites) return fail(NLS_ERR_ARG, "nls_reduce_mass_noalloc: traj_stride < n_sites");
  CU(cudaSetDevice(device));
  unsigned char *scratch;
  int nsm;
  {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) { // set-up, once per device, with that device current
      unsigned char *p = nullptr;
      CU(cudaGetSymbolAddress((void **)&p, g_noalloc_scratch));
      if (!p) return fail(NLS_ERR_CUDA, "nls_reduce_mass_noalloc: cudaGetSymbolAddress returned a null address");
      int n = 0;
      CU(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device));
      g_noalloc_nsm[device] = n > 0 ? n : 148;
      g_noalloc_ptr[device] = p;
    }
    scratch = g_noalloc_ptr[device];
    nsm = g_noalloc_nsm[device];
  }
  RedCtx rc;
  rc.partials = (double *)scratch;
  rc.tickets = (unsigned *)(scratch + (size_t)kNaCtas * 8);
  rc.slab = nullptr;
  // CTAs per trajectory: one wave of 8 CTAs/SM over the batch, within the scratch's kNaCtas partial slots
  long long per = (8LL * nsm) / batch;
  const long long want = (long long)((n_sites + kThreads - 1) / kThreads);
  if (per > want) per = want;
  if (per > (long long)(kNaCtas / batch)) per = kNaCtas / batch;
  if (per < 1) per = 1;
  cudaStream_t st = (cudaStream_t)stream;
  // the tickets must start at zero (the array keeps whatever the last -- possibly aborted -- call left there)
  CU(cudaMemsetAsync(rc.tickets, 0, (size_t)batch * sizeof(unsigned), st));
  const dim3 grid((unsigned)per, batch);
  if (is_complex)
    k_reduce_mass_noalloc<double2><<<grid, kThreads, 0, st>>>((const double2 *)dev_u, (long long)n_sites, (long long)traj_stride, cell, dev_out, rc);
  else
    k_reduce_mass_noalloc<double><<<grid, kThreads, 0, st>>>((const double *)dev_u, (long long)n_sites, (long long)traj_stride, cell, dev_out, rc);
  CU(cudaGetLastError());
  return NLS_OK;
}

int nls_slab_export(nls_handle e, void *buf128) {
  if (!e || !buf128) return fail(NLS_ERR_ARG, "null argument");
  if (!e->slab) return fail(NLS_ERR_STATE, "not a z-slab handle (nz_global == nz)");
  CU(cudaSetDevice(e->d.device));
  SlabDesc sd;
  std::memset(&sd, 0, sizeof sd);
  CU(cudaIpcGetMemHandle(&sd.ipc, e->arena));
  sd.magic = kSlabMagic;
  sd.pid = (uint64_t)getpid();
  sd.raw = (uint64_t)(uintptr_t)e->arena;
  sd.stride = (uint64_t)e->g.stride;
  sd.nz = e->d.nz; sd.z0 = e->d.z0; sd.nx = e->d.nx; sd.ny = e->d.ny;
  sd.device = e->d.device; sd.esz = (int32_t)e->esz;
  std::memcpy(buf128, &sd, sizeof sd);
  return NLS_OK;
}

int nls_slab_connect(nls_handle e, const void *descs, int rank, int nranks) {
  if (!e || !descs) return fail(NLS_ERR_ARG, "null argument");
  if (!e->slab) return fail(NLS_ERR_STATE, "not a z-slab handle (nz_global == nz)");
  if (e->connected) return fail(NLS_ERR_STATE, "already connected");
  if (nranks < 2 || nranks > kSlabMaxRanks || rank < 0 || rank >= nranks)
    return fail(NLS_ERR_ARG, "need 2..8 ranks and 0 <= rank < nranks");
  CU(cudaSetDevice(e->d.device));
  const SlabDesc *sd = (const SlabDesc *)descs;
  uint32_t z = 0;
  for (int q = 0; q < nranks; ++q) { // slabs must tile [0, nz_global) in rank order
    if (sd[q].magic != kSlabMagic) return fail(NLS_ERR_ARG, "bad slab descriptor");
    if (sd[q].nx != e->d.nx || sd[q].ny != e->d.ny || sd[q].esz != (int32_t)e->esz || sd[q].z0 != z)
      return fail(NLS_ERR_ARG, "slab descriptors do not tile the global grid in rank order");
    z += sd[q].nz;
  }
  if (z != (uint32_t)e->g.nz_glob || sd[rank].z0 != e->d.z0 || sd[rank].nz != e->d.nz ||
      sd[rank].raw != (uint64_t)(uintptr_t)e->arena || sd[rank].pid != (uint64_t)getpid())
    return fail(NLS_ERR_ARG, "descriptor[rank] is not this handle / slabs do not cover nz_global");
  SlabCtx sc;
  std::memset(&sc, 0, sizeof sc);
  sc.rank = rank; sc.nranks = nranks; sc.status = e->status;
  for (int q = 0; q < nranks; ++q) {
    nls_engine::Peer &p = e->peers[q];
    p.stride = sd[q].stride; p.nz = sd[q].nz;
    if (q == rank) p.base = e->arena;
    else if (sd[q].pid == (uint64_t)getpid()) { // same process (one host thread per GPU): direct peer access
      if (sd[q].device != e->d.device) {
        int can = 0;
        CU(cudaDeviceCanAccessPeer(&can, e->d.device, sd[q].device));
        if (!can) return fail(NLS_ERR_CUDA, "no peer access between the slab devices");
        cudaError_t pe = cudaDeviceEnablePeerAccess(sd[q].device, 0);
        if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled)
          return fail(NLS_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(pe));
        cudaGetLastError();
      }
      p.base = (char *)(uintptr_t)sd[q].raw;
    } else {
      void *m = nullptr;
      CU(cudaIpcOpenMemHandle(&m, sd[q].ipc, cudaIpcMemLazyEnablePeerAccess));
      p.base = (char *)m; p.ipc = true;
    }
    sc.box[q] = (SlabMailbox *)p.base;
  }
  CU(cudaMemcpy(e->slab_dev, &sc, sizeof sc, cudaMemcpyHostToDevice));
  e->rank = rank; e->nranks = nranks;
  e->rc.slab = e->slab_dev;
  e->connected = true;
  e->start_valid = false;
  return NLS_OK;
}

} // extern "C"
