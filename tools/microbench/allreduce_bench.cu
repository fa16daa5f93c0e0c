// allreduce_bench.cu -- micro-benchmark of grid-wide "barrier + all-reduce of K doubles" schemes for the persistent
// cooperative kernels (nls_persist.cuh): 148 CTAs x 512 threads, R rounds, optional per-round global stores (the
// phase's output that the release has to drain).  Prints microseconds per round for every scheme.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o allreduce_bench allreduce_bench.cu && ./allreduce_bench
#include <cooperative_groups.h>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

constexpr int kThreads = 512, kWarps = 16, kStride = 160, kMaxK = 40;
constexpr unsigned long long kSent = 0xffffffffffffffffull;

__device__ __forceinline__ unsigned ld_acq_u32(const unsigned *p) { unsigned v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ unsigned ld_rlx_u32(const unsigned *p) { unsigned v; asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void red_rel_add(unsigned *p, unsigned v) { asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned long long ld_acq_u64(const unsigned long long *p) { unsigned long long v; asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ unsigned long long ld_rlx_u64(const unsigned long long *p) { unsigned long long v; asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_rel_u64(unsigned long long *p, unsigned long long v) { asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ void st_rlx_u64(unsigned long long *p, unsigned long long v) { asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ void fence_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

struct Args {
  int K, R, scheme, work;    // work: doubles stored per thread per round before the reduction
  double *part;              // [3][kMaxK][kStride*4]   (x4: room for the padded layouts)
  unsigned *ctr;             // [64]
  double *tot;               // [3][kMaxK + 8]  totals + sequence word (two-level scheme)
  double *sink;              // work target
  double *out;
};

__device__ __forceinline__ void cta_reduce(int K, const double *acc, double (*s_red)[kWarps], double *s_val) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int k = 0; k < K; ++k) {
    double v = acc[k];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) s_red[k][warp] = v;
  }
  __syncthreads();
  if (threadIdx.x < K) {
    double v = 0.0;
    for (int w = 0; w < kWarps; ++w) v += s_red[threadIdx.x][w];
    s_val[threadIdx.x] = v;
  }
}

// gather helpers: sum of `ncta` entries at p[c*estride], all loads of a lane in flight together
__device__ __forceinline__ double warp_sum_plain(const double *p, int ncta, int estride) {
  const int lane = threadIdx.x & 31;
  double t[5];
  for (int q = 0; q < 5; ++q) { const int c = lane + 32 * q; t[q] = c < ncta ? __ldcg(p + (size_t)c * estride) : 0.0; }
  double v = 0.0;
  for (int q = 0; q < 5; ++q) v += t[q];
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
template <bool ACQ, int SLEEP>
__device__ __forceinline__ double warp_poll_sum(const double *p, int ncta, int estride) {
  const int lane = threadIdx.x & 31;
  unsigned long long t[5];
  for (int q = 0; q < 5; ++q) t[q] = lane + 32 * q < ncta ? kSent : 0ull;
  for (;;) {
    bool pend = false;
    for (int q = 0; q < 5; ++q)
      if (t[q] == kSent) {
        const unsigned long long *a = (const unsigned long long *)p + (size_t)(lane + 32 * q) * estride;
        t[q] = ACQ ? ld_acq_u64(a) : ld_rlx_u64(a);
      }
    for (int q = 0; q < 5; ++q) pend |= t[q] == kSent;
    if (!__any_sync(0xffffffffu, pend)) break;
    if (SLEEP) __nanosleep(SLEEP);
  }
  if (!ACQ) fence_gpu();
  double v = 0.0;
  for (int q = 0; q < 5; ++q) v += __longlong_as_double((long long)t[q]);
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(kThreads, 1) k_bench(Args a) {
  __shared__ double s_red[kMaxK][kWarps];
  __shared__ double s_val[kMaxK], s_out[kMaxK];
  const int K = a.K, ncta = gridDim.x, cta = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double acc[kMaxK];
  for (int k = 0; k < kMaxK; ++k) acc[k] = 1e-3 * (threadIdx.x + 1) + k + 1e-6 * cta;
  unsigned target = 0;
  double check = 0.0;
  cg::grid_group grid = cg::this_grid();
  for (int r = 0; r < a.R; ++r) {
    // the phase's output
    for (int q = 0; q < a.work; ++q) a.sink[((size_t)cta * a.work + q) * kThreads + threadIdx.x] = acc[0] + r + q;
    cta_reduce(K, acc, s_red, s_val);
    const int slot2 = r & 1, slot3 = r % 3;
    if (a.scheme == 0) { // round-1 scheme: partials, counter barrier (thread 0), gather
      double *pp = a.part + (size_t)slot2 * kMaxK * kStride;
      if (threadIdx.x < K) pp[(size_t)threadIdx.x * kStride + cta] = s_val[threadIdx.x];
      __syncthreads();
      if (threadIdx.x == 0) {
        target += ncta;
        red_rel_add(a.ctr, 1u);
        while (ld_acq_u32(a.ctr) < target) {}
      }
      __syncthreads();
      for (int k = warp; k < K; k += kWarps) { const double v = warp_sum_plain(pp + (size_t)k * kStride, ncta, 1); if (lane == 0) s_out[k] = v; }
      __syncthreads();
    } else if (a.scheme == 1 || a.scheme == 2 || a.scheme == 3 || a.scheme == 4 || a.scheme == 6 || a.scheme == 7) {
      // sentinel polling. 1: warps poll (acquire); 2: every warp polls every k; 3: relaxed + fence; 4: padded to 32-byte
      // sectors; 6: nanosleep back-off; 7: 128-byte padding
      const int es = a.scheme == 4 ? 4 : a.scheme == 7 ? 16 : 1;
      unsigned long long *pp = (unsigned long long *)a.part + (size_t)slot3 * kMaxK * kStride * 16;
      unsigned long long *pn = (unsigned long long *)a.part + (size_t)((slot3 + 1) % 3) * kMaxK * kStride * 16;
      if (threadIdx.x < K) st_rlx_u64(pn + ((size_t)threadIdx.x * kStride + cta) * es, kSent);
      __syncthreads();
      if (threadIdx.x < K) st_rel_u64(pp + ((size_t)threadIdx.x * kStride + cta) * es, (unsigned long long)__double_as_longlong(s_val[threadIdx.x]));
      if (a.scheme == 2) {
        for (int k = 0; k < K; ++k) { const double v = warp_poll_sum<true, 0>((const double *)pp + (size_t)k * kStride * es, ncta, es); if (lane == 0 && warp == k % kWarps) s_out[k] = v; }
      } else {
        for (int k = warp; k < K; k += kWarps) {
          double v;
          if (a.scheme == 3) v = warp_poll_sum<false, 0>((const double *)pp + (size_t)k * kStride * es, ncta, es);
          else if (a.scheme == 6) v = warp_poll_sum<true, 40>((const double *)pp + (size_t)k * kStride * es, ncta, es);
          else v = warp_poll_sum<true, 0>((const double *)pp + (size_t)k * kStride * es, ncta, es);
          if (lane == 0) s_out[k] = v;
        }
      }
      __syncthreads();
    } else if (a.scheme == 5) { // two-level: CTA 0 gathers the partials (sentinel polling), publishes totals + sequence word
      unsigned long long *pp = (unsigned long long *)a.part + (size_t)slot3 * kMaxK * kStride * 16;
      unsigned long long *pn = (unsigned long long *)a.part + (size_t)((slot3 + 1) % 3) * kMaxK * kStride * 16;
      unsigned long long *tt = (unsigned long long *)a.tot + (size_t)slot3 * (kMaxK + 8);
      if (threadIdx.x < K) st_rlx_u64(pn + (size_t)threadIdx.x * kStride + cta, kSent);
      __syncthreads();
      if (threadIdx.x < K) st_rel_u64(pp + (size_t)threadIdx.x * kStride + cta, (unsigned long long)__double_as_longlong(s_val[threadIdx.x]));
      if (cta == 0) {
        for (int k = warp; k < K; k += kWarps) { const double v = warp_poll_sum<true, 0>((const double *)pp + (size_t)k * kStride, ncta, 1); if (lane == 0) s_out[k] = v; }
        __syncthreads();
        if (threadIdx.x < K) st_rlx_u64(tt + threadIdx.x, (unsigned long long)__double_as_longlong(s_out[threadIdx.x]));
        __syncthreads();
        if (threadIdx.x == 0) st_rel_u64(tt + kMaxK, (unsigned long long)(r + 1));
      } else {
        if (threadIdx.x == 0) { while (ld_acq_u64(tt + kMaxK) != (unsigned long long)(r + 1)) {} }
        __syncthreads();
        if (threadIdx.x < K) s_out[threadIdx.x] = __longlong_as_double((long long)ld_rlx_u64(tt + threadIdx.x));
      }
      __syncthreads();
    } else if (a.scheme == 8) { // cooperative groups grid.sync + gather (library baseline)
      double *pp = a.part + (size_t)slot2 * kMaxK * kStride;
      if (threadIdx.x < K) pp[(size_t)threadIdx.x * kStride + cta] = s_val[threadIdx.x];
      grid.sync();
      for (int k = warp; k < K; k += kWarps) { const double v = warp_sum_plain(pp + (size_t)k * kStride, ncta, 1); if (lane == 0) s_out[k] = v; }
      __syncthreads();
    } else if (a.scheme == 9) { // counter barrier with the arrival spread over 8 counters (hash by cta), thread 0 sums them
      double *pp = a.part + (size_t)slot2 * kMaxK * kStride;
      if (threadIdx.x < K) pp[(size_t)threadIdx.x * kStride + cta] = s_val[threadIdx.x];
      __syncthreads();
      if (warp == 0) {
        if (lane == 0) red_rel_add(a.ctr + 32 * (cta & 7), 1u);
        target += ncta;
        for (;;) {
          unsigned v = lane < 8 ? ld_acq_u32(a.ctr + 32 * lane) : 0u;
          for (int o = 4; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
          if (__shfl_sync(0xffffffffu, v, 0) >= target) break;
        }
      }
      __syncthreads();
      for (int k = warp; k < K; k += kWarps) { const double v = warp_sum_plain(pp + (size_t)k * kStride, ncta, 1); if (lane == 0) s_out[k] = v; }
      __syncthreads();
    }
    check += s_out[0] + (K > 1 ? s_out[K - 1] : 0.0);
    for (int k = 0; k < K; ++k) acc[k] += 1e-9 * s_out[k % K]; // dependence on the result
  }
  if (threadIdx.x == 0) a.out[cta] = check;
}

int main(int argc, char **argv) {
  int dev = 0, nsm = 0;
  cudaSetDevice(dev);
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  Args a;
  const size_t pbytes = (size_t)3 * kMaxK * kStride * 16 * 8;
  cudaMalloc(&a.part, pbytes);
  cudaMalloc(&a.ctr, 64 * 32 * 4);
  cudaMalloc(&a.tot, 3 * (kMaxK + 8) * 8);
  cudaMalloc(&a.sink, (size_t)nsm * 64 * kThreads * 8);
  cudaMalloc(&a.out, nsm * 8);
  a.R = 2000;
  const int works[] = {0, 14};
  const int Ks[] = {1, 10, 20};
  const int schemes[] = {0, 8, 9, 1, 2, 3, 4, 7, 6, 5};
  const char *names[] = {"0 counter barrier + gather (round 1)", "1 sentinel poll, warp per k (acquire)", "2 sentinel poll, every warp polls", "3 sentinel poll, relaxed + fence",
                         "4 sentinel poll, 32 B padding", "5 two-level (CTA 0 reduces, totals + seq)", "6 sentinel poll + nanosleep(40)", "7 sentinel poll, 128 B padding",
                         "8 cg::grid.sync + gather", "9 counter barrier over 8 counters + gather"};
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int w : works)
    for (int K : Ks) {
      printf("---- K = %d doubles, %d CTAs x %d threads, %d B stored per CTA per round before the reduction\n", K, nsm, kThreads, w * kThreads * 8);
      for (int s : schemes) {
        a.K = K; a.scheme = s; a.work = w;
        double best = 1e30, chk = 0;
        for (int rep = 0; rep < 3; ++rep) {
          cudaMemset(a.part, 0xff, pbytes);
          cudaMemset(a.ctr, 0, 64 * 32 * 4);
          cudaMemset(a.tot, 0, 3 * (kMaxK + 8) * 8);
          void *args[] = {&a};
          cudaEventRecord(e0);
          cudaError_t ce = cudaLaunchCooperativeKernel((const void *)k_bench, dim3(nsm), dim3(kThreads), args, 0, 0);
          cudaEventRecord(e1);
          cudaEventSynchronize(e1);
          if (ce != cudaSuccess || cudaGetLastError() != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(ce)); return 1; }
          float ms;
          cudaEventElapsedTime(&ms, e0, e1);
          if (ms < best) best = ms;
          cudaMemcpy(&chk, a.out, 8, cudaMemcpyDeviceToHost);
        }
        printf("  scheme %-46s %7.3f us/round   (check %.6e)\n", names[s], 1e3 * best / a.R, chk);
      }
    }
  return 0;
}
