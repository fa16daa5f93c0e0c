// tridiag_bench.cu -- accuracy + latency of the in-kernel m x m tridiagonal "f(T) e1" solvers on realistic Lanczos
// matrices (2D no-flux Laplacian, white-noise / smooth / nearly invariant start vectors, m = 10..30, T(m-1,m-1) = 0
// as in the reference).  Variants:
//   reg    register-resident warp QL (tridiag_warp_reg, nls_kernels.cuh)
//   bis    lane-parallel: eigenvalue k by Sturm bisection in lane k, eigenvector by twisted factorisation
// Reference: cyclic Jacobi in long double on the host.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I ../../include -o tridiag_bench tridiag_bench.cu
#include "../../nonlinear-solvers_b200/csrc/nls_kernels.cuh"
#include <cmath>
#include <complex>
#include <cstdio>
#include <random>
#include <vector>
using namespace nls;

// ---------------------------------------------------------------- lane-parallel bisection + twisted factorisation
// d, e: lane-replicated copies in shared memory ([i][32] layout not needed: read-only, broadcast reads)
__device__ __forceinline__ int sturm_lt(const double *d, const double *e2, int m, double x) { // #eigenvalues < x
  double pp = 1.0, pc = d[0] - x;
  int cnt = pc < 0.0;
  for (int i = 1; i < m; ++i) {
    double pn = fma(d[i] - x, pc, -e2[i - 1] * pp);
    if (pn == 0.0) pn = pc < 0 ? 1e-300 : -1e-300; // an exact zero counts as a sign change
    cnt += (pn < 0.0) != (pc < 0.0);
    pp = pc; pc = pn;
    if (fabs(pc) > 1e150) { pc *= 1e-150; pp *= 1e-150; }
    else if (fabs(pc) < 1e-150) { pc *= 1e150; pp *= 1e150; }
  }
  return cnt;
}
// out: lam (this lane's eigenvalue, lane < m), z column in zs[i*32 + lane] (normalised)
__device__ void tridiag_bisect_warp(const double *d, const double *e, const double *e2, int m, double *work /*[2][m][32]*/,
                                    double *zs /*[m][32]*/, double &lam) {
  const int lane = threadIdx.x & 31;
  double lo = 1e300, hi = -1e300, tn = 0.0;
  for (int i = 0; i < m; ++i) {
    const double r = (i > 0 ? fabs(e[i - 1]) : 0.0) + (i < m - 1 ? fabs(e[i]) : 0.0);
    lo = fmin(lo, d[i] - r); hi = fmax(hi, d[i] + r);
    tn = fmax(tn, fabs(d[i]) + r);
  }
  lo -= 1e-14 * tn; hi += 1e-14 * tn;
  if (lane < m) {
    const int k = lane; // k-th smallest
    for (int it = 0; it < 60; ++it) {
      const double mid = 0.5 * (lo + hi);
      if (mid <= lo || mid >= hi) break;
      if (sturm_lt(d, e2, m, mid) > k) hi = mid; else lo = mid;
    }
    lam = 0.5 * (lo + hi);
    // twisted factorisation (LAPACK dlar1v): forward / backward pivots, twist at the smallest |gamma|
    double *Dp = work + lane, *Dm = work + m * 32 + lane;
    const double piv = 2.3e-16 * tn + 1e-300;
    double t = d[0] - lam;
    for (int i = 0; i < m - 1; ++i) {
      if (fabs(t) < piv) t = t < 0 ? -piv : piv;
      Dp[i * 32] = t;
      t = (d[i + 1] - lam) - e2[i] / t;
    }
    if (fabs(t) < piv) t = t < 0 ? -piv : piv;
    Dp[(m - 1) * 32] = t;
    t = d[m - 1] - lam;
    for (int i = m - 1; i > 0; --i) {
      if (fabs(t) < piv) t = t < 0 ? -piv : piv;
      Dm[i * 32] = t;
      t = (d[i - 1] - lam) - e2[i - 1] / t;
    }
    if (fabs(t) < piv) t = t < 0 ? -piv : piv;
    Dm[0] = t;
    int r = 0;
    double gmin = 1e300;
    for (int i = 0; i < m; ++i) {
      const double g = fabs(Dp[i * 32] + Dm[i * 32] - (d[i] - lam));
      if (g < gmin) { gmin = g; r = i; }
    }
    double *z = zs + lane;
    z[r * 32] = 1.0;
    double nrm = 1.0;
    for (int i = r - 1; i >= 0; --i) { const double v = -(e[i] / Dp[i * 32]) * z[(i + 1) * 32]; z[i * 32] = v; nrm = fma(v, v, nrm); }
    for (int i = r; i < m - 1; ++i) { const double v = -(e[i] / Dm[(i + 1) * 32]) * z[i * 32]; z[(i + 1) * 32] = v; nrm = fma(v, v, nrm); }
    const double inv = rsqrt(nrm);
    for (int i = 0; i < m; ++i) z[i * 32] *= inv;
  }
  __syncwarp();
}

struct Case { int m; double d[32], e[32]; double tr, ti; int fn; };

__global__ void k_solve(const Case *cs, int ncase, int variant, int reps, double *out /*[ncase][64]*/, long long *cycles) {
  __shared__ double Z[33 * 33], work[2 * 32 * 32], zs[32 * 32], zf[64], sd[32], se[32], se2[32], co[128];
  const int lane = threadIdx.x;
  for (int c = 0; c < ncase; ++c) {
    const Case &cc = cs[c];
    const int m = cc.m;
    TridiagArgs ta;
    ta.m = m; ta.W = 1; ta.nfn = 1; ta.fn[0] = cc.fn; ta.tr[0] = cc.tr; ta.ti[0] = cc.ti;
    long long t0 = clock64();
    for (int rep = 0; rep < reps; ++rep) {
      if (variant == 0) {
        double dr = lane < m ? cc.d[lane] : 0.0, er = lane < m - 1 ? cc.e[lane] : 0.0;
        if (lane < m) for (int k = 0; k < m; ++k) Z[lane * (m + 1) + k] = lane == k ? 1.0 : 0.0;
        __syncwarp();
        tridiag_warp_reg(dr, er, Z, m + 1, m);
        __syncwarp();
        tridiag_coefs_reg(dr, Z, m + 1, m, ta, 1.0, zf, co, 128);
      } else {
        if (lane < m) { sd[lane] = cc.d[lane]; se[lane] = lane < m - 1 ? cc.e[lane] : 0.0; se2[lane] = se[lane] * se[lane]; }
        __syncwarp();
        double lam = 0.0;
        tridiag_bisect_warp(sd, se, se2, m, work, zs, lam);
        // co_i = sum_k z_i^(k) z_0^(k) f(lam_k)
        if (lane < m) { const double2 fv = fn_eval(cc.fn, cc.tr, cc.ti, lam); const double z0 = zs[lane]; zf[2 * lane] = fv.x * z0; zf[2 * lane + 1] = fv.y * z0; }
        __syncwarp();
        if (lane < m) {
          double sr = 0, si = 0;
          for (int k = 0; k < m; ++k) { const double z = zs[lane * 32 + k]; sr = fma(z, zf[2 * k], sr); si = fma(z, zf[2 * k + 1], si); }
          co[2 * lane] = sr; co[2 * lane + 1] = si;
        }
        __syncwarp();
      }
    }
    long long t1 = clock64();
    if (lane == 0) cycles[c] = (t1 - t0) / reps;
    if (lane < m) { out[c * 64 + 2 * lane] = co[2 * lane]; out[c * 64 + 2 * lane + 1] = co[2 * lane + 1]; }
    __syncwarp();
  }
}

// ---------------------------------------------------------------- host: Lanczos matrices + long double reference
typedef long double ld;
static void jacobi_eig(int m, std::vector<ld> &A, std::vector<ld> &V) { // cyclic Jacobi, A symmetric m x m (row-major)
  V.assign(m * m, 0); for (int i = 0; i < m; ++i) V[i * m + i] = 1;
  for (int sweep = 0; sweep < 60; ++sweep) {
    ld off = 0; for (int i = 0; i < m; ++i) for (int j = i + 1; j < m; ++j) off += A[i * m + j] * A[i * m + j];
    if (off < 1e-60L) break;
    for (int p = 0; p < m; ++p) for (int q = p + 1; q < m; ++q) {
      if (fabsl(A[p * m + q]) < 1e-4000L) continue;
      ld th = (A[q * m + q] - A[p * m + p]) / (2 * A[p * m + q]);
      ld t = (th >= 0 ? 1 : -1) / (fabsl(th) + sqrtl(th * th + 1));
      ld c = 1 / sqrtl(t * t + 1), s = t * c;
      for (int k = 0; k < m; ++k) { ld akp = A[k * m + p], akq = A[k * m + q]; A[k * m + p] = c * akp - s * akq; A[k * m + q] = s * akp + c * akq; }
      for (int k = 0; k < m; ++k) { ld apk = A[p * m + k], aqk = A[q * m + k]; A[p * m + k] = c * apk - s * aqk; A[q * m + k] = s * apk + c * aqk; }
      for (int k = 0; k < m; ++k) { ld vkp = V[k * m + p], vkq = V[k * m + q]; V[k * m + p] = c * vkp - s * vkq; V[k * m + q] = s * vkp + c * vkq; }
    }
  }
}
static std::complex<ld> f_host(int fn, ld tr, ld ti, ld lam) {
  if (fn == NLS_FN_EXP) return std::exp(std::complex<ld>(tr * lam, ti * lam));
  if (fn == NLS_FN_COS_SQRT) return cosl(tr * sqrtl(fabsl(lam)));
  ld x = tr / 2 * sqrtl(fabsl(lam)); ld s = fabsl(x) < 1e-8L ? 1 : sinl(x) / x; return s * s; // SINC2_HALF
}
static void lanczos2d(int nx, double dx, const std::vector<double> &u, int m, double *d, double *e) { // real, full MGS
  const int n = nx * nx;
  auto L = [&](const std::vector<double> &x, std::vector<double> &y) {
    for (int j = 0; j < nx; ++j) for (int i = 0; i < nx; ++i) {
      const int p = j * nx + i; const bool ring = i == 0 || i == nx - 1 || j == 0 || j == nx - 1;
      double a = (ring ? -3.0 : -4.0) * x[p];
      if (i > 0) a += x[p - 1]; if (i < nx - 1) a += x[p + 1]; if (j > 0) a += x[p - nx]; if (j < nx - 1) a += x[p + nx];
      y[p] = a / (dx * dx);
    }
  };
  std::vector<std::vector<double>> V(m, std::vector<double>(n));
  double b = 0; for (double v : u) b += v * v; b = std::sqrt(b);
  for (int p = 0; p < n; ++p) V[0][p] = u[p] / b;
  std::vector<double> w(n);
  for (int j = 0; j < m; ++j) { d[j] = 0; e[j] = 0; }
  for (int j = 0; j < m - 1; ++j) {
    L(V[j], w);
    if (j > 0) for (int p = 0; p < n; ++p) w[p] -= e[j - 1] * V[j - 1][p];
    double a = 0; for (int p = 0; p < n; ++p) a += V[j][p] * w[p];
    d[j] = a;
    for (int p = 0; p < n; ++p) w[p] -= a * V[j][p];
    for (int i = 0; i <= j; ++i) { double h = 0; for (int p = 0; p < n; ++p) h += V[i][p] * w[p]; for (int p = 0; p < n; ++p) w[p] -= h * V[i][p]; }
    double nb = 0; for (double v : w) nb += v * v; nb = std::sqrt(nb);
    e[j] = nb;
    for (int p = 0; p < n; ++p) V[j + 1][p] = w[p] / nb;
  }
}

int main() {
  std::vector<Case> cases;
  std::vector<const char *> names;
  std::mt19937 rng(42);
  std::normal_distribution<double> nd(0, 1);
  const int nx = 96; const double Lb = 10.0, dx = 2 * Lb / (nx - 1);
  for (int kind = 0; kind < 3; ++kind)
    for (int m : {10, 12, 20, 30}) {
      std::vector<double> u(nx * nx);
      for (int j = 0; j < nx; ++j) for (int i = 0; i < nx; ++i) {
        const double x = -Lb + i * dx, y = -Lb + j * dx;
        if (kind == 0) u[j * nx + i] = nd(rng);
        else if (kind == 1) u[j * nx + i] = std::exp(-(x * x + y * y) / 2) * std::cos(x) + 0.3 * std::exp(-((x - 3) * (x - 3) + (y + 2) * (y + 2)));
        else u[j * nx + i] = std::cos(M_PI * (i + 0.5) / nx) * std::cos(2 * M_PI * (j + 0.5) / nx) + 1e-9 * nd(rng); // ~ eigenvector: Krylov space exhausted at once
      }
      for (int fn : {NLS_FN_EXP, NLS_FN_COS_SQRT, NLS_FN_SINC2_HALF}) {
        Case c; c.m = m; lanczos2d(nx, dx, u, m, c.d, c.e);
        c.fn = fn; c.tr = fn == NLS_FN_EXP ? 0.0 : 1e-3; c.ti = fn == NLS_FN_EXP ? 1e-3 : 0.0;
        cases.push_back(c);
        names.push_back(kind == 0 ? "noise" : kind == 1 ? "smooth" : "near-eigvec");
      }
    }
  const int nc = (int)cases.size();
  Case *dc; double *dout; long long *dcy;
  cudaMalloc(&dc, nc * sizeof(Case)); cudaMalloc(&dout, nc * 64 * 8); cudaMalloc(&dcy, nc * 8);
  cudaMemcpy(dc, cases.data(), nc * sizeof(Case), cudaMemcpyHostToDevice);
  std::vector<std::vector<double>> res(2, std::vector<double>(nc * 64));
  std::vector<std::vector<long long>> cyc(2, std::vector<long long>(nc));
  for (int v = 0; v < 2; ++v) {
    k_solve<<<1, 32>>>(dc, nc, v, 20, dout, dcy);
    cudaError_t ce = cudaDeviceSynchronize();
    if (ce != cudaSuccess) { printf("variant %d failed: %s\n", v, cudaGetErrorString(ce)); return 1; }
    cudaMemcpy(res[v].data(), dout, nc * 64 * 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(cyc[v].data(), dcy, nc * 8, cudaMemcpyDeviceToHost);
  }
  printf("%-12s %3s %3s | %10s %10s | %12s %12s | min gap/||T||  beta_min/||T||\n", "start", "m", "fn", "cyc reg-QL", "cyc bisect", "err reg-QL", "err bisect");
  for (int c = 0; c < nc; ++c) {
    const Case &cc = cases[c]; const int m = cc.m;
    std::vector<ld> A(m * m, 0), V;
    for (int i = 0; i < m; ++i) { A[i * m + i] = cc.d[i]; if (i < m - 1) A[i * m + i + 1] = A[(i + 1) * m + i] = cc.e[i]; }
    jacobi_eig(m, A, V);
    ld tn = 0, gap = 1e300L, bmin = 1e300L;
    for (int i = 0; i < m; ++i) tn = std::max(tn, fabsl(A[i * m + i]));
    for (int i = 0; i < m; ++i) for (int j = i + 1; j < m; ++j) gap = std::min(gap, fabsl(A[i * m + i] - A[j * m + j]));
    for (int i = 0; i < m - 1; ++i) bmin = std::min(bmin, (ld)fabs(cc.e[i]));
    std::vector<std::complex<ld>> ref(m);
    for (int i = 0; i < m; ++i) { std::complex<ld> s = 0; for (int k = 0; k < m; ++k) s += V[i * m + k] * V[0 * m + k] * f_host(cc.fn, cc.tr, cc.ti, A[k * m + k]); ref[i] = s; }
    double err[2];
    for (int v = 0; v < 2; ++v) {
      ld num = 0, den = 0;
      for (int i = 0; i < m; ++i) { std::complex<ld> g(res[v][c * 64 + 2 * i], res[v][c * 64 + 2 * i + 1]); num += std::norm(g - ref[i]); den += std::norm(ref[i]); }
      err[v] = (double)sqrtl(num / den);
    }
    printf("%-12s %3d %3d | %10lld %10lld | %12.3e %12.3e | %10.2e %10.2e\n", names[c], m, cc.fn, cyc[0][c], cyc[1][c], err[0], err[1], (double)(gap / tn), (double)(bmin / tn));
  }
  return 0;
}
