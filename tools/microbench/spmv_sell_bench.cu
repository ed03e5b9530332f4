// Microbenchmark: direct CSR kernel (spmv_direct.cuh) against the sliced-ELL kernel
// (spmv_sell.cuh) on synthetic matrices of BASELINE config 2's shapes:
//   A-like: m = 1e6 rows x n = 1e5 columns, exactly 10 random columns per row
//   M-like: n = 1e5 rows, 7 entries with col < n and 100 +- 10 entries with col in [n, n + m)
// Build variants with -DSELL_TMA=0/1 -DSELL_CH=.. -DSELL_NST=.. -DSELL_MINB=.. -DSELL_LGM=..
#include "../../osqp_b200/csrc/spmv_sell.cuh"
#include "../../osqp_b200/csrc/spmv_direct.cuh"
#include <algorithm>
#include <cmath>
#ifndef SELL_LGM
#define SELL_LGM 2
#endif
int dev_cache_alloc(void **p, size_t bytes) { return cudaMalloc(p, bytes) != cudaSuccess; }
void dev_cache_free(void *p) { cudaFree(p); }

struct EpiScale {   // w = rho .* sum, the shape of the CG operator's first half
  static constexpr int NSUM = 1, NMAX = 0;
  const double *rho; double *w;
  __device__ bool begin(double *) { return true; }
  struct Pre { double rho; };
  __device__ __forceinline__ Pre load(int i) const { return Pre{rho[i]}; }
  __device__ void row(int i, const double *s, const Pre &pre, double *asum, double *) const {
    w[i] = pre.rho * s[0]; asum[0] += s[0];
  }
};
static unsigned long long rs = 88172645463325252ull;
static inline unsigned long long rnd() { rs ^= rs << 13; rs ^= rs >> 7; rs ^= rs << 17; return rs; }

struct Host { int rows, cols; std::vector<int> rp, col; std::vector<double> val; };
template <class T> T *up(const std::vector<T> &h, size_t extra = 8) {
  T *d; cudaMalloc(&d, (h.size() + extra) * sizeof(T)); cudaMemset(d, 0, (h.size() + extra) * sizeof(T));
  cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice); return d;
}
template <class F> float timeit(F f, int reps = 20) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int r = 0; r < 3; ++r) f();
  cudaEventRecord(a);
  for (int r = 0; r < reps; ++r) f();
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms / reps * 1000.f;
}

static void run(const char *name, const Host &H, int lg, int split_col, int nlo, int nhi) {
  const int rows = H.rows, nnz = (int)H.col.size();
  std::vector<double> xlo(nlo), xhi(nhi > 0 ? nhi : 1), rho(rows);
  for (auto &t : xlo) t = (double)(rnd() % 2001) / 1000.0 - 1.0;
  for (auto &t : xhi) t = (double)(rnd() % 2001) / 1000.0 - 1.0;
  for (auto &t : rho) t = 0.5 + (double)(rnd() % 1000) / 1000.0;
  double *dxlo = up(xlo), *dxhi = up(xhi), *drho = up(rho), *w0, *w1, *part;
  cudaMalloc(&w0, rows * 8); cudaMalloc(&w1, rows * 8); cudaMalloc(&part, 8 * 4096);
  // ---- CSR + direct kernel
  CsrView C{};
  C.rows = rows; C.nnz = nnz; C.group = spmv_direct_group(nnz, rows); C.split_col = split_col;
  C.rowptr = up(H.rp); C.col = up(H.col); C.val = up(H.val);
  Gather g0{dxlo, dxhi}, g1{dxlo, dxhi};
  EpiScale e0{drho, w0}, e1{drho, w1};
  const int gridD = OSQP_CUDA_SMS * SPMV_DIRECT_MINB;
  cudaFuncSetAttribute(spmv_direct_kernel<1, false, 0, EpiScale>,
                       cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxL1);
  const float td = timeit([&] { spmv_direct_kernel<1, false, 0, EpiScale><<<gridD, OSQP_CUDA_THREADS>>>(C, g0, g1, e0, part); });
  // ---- SELL on the host
  const int G = 1 << lg, rps = 32 >> lg, ns = (rows + rps - 1) / rps;
  std::vector<int> sp(ns + 1, 0);
  for (int s = 0; s < ns; ++s) {
    int k = 0;
    for (int r = s * rps; r < std::min(rows, (s + 1) * rps); ++r) k = std::max(k, (H.rp[r + 1] - H.rp[r] + G - 1) / G);
    sp[s + 1] = sp[s] + k;
  }
  const int tp = sp[ns];
  std::vector<int> sc((size_t)tp * 32, -1); std::vector<double> sv((size_t)tp * 32, 0.0);
  for (int s = 0; s < ns; ++s)
    for (int p = 0; p < sp[s + 1] - sp[s]; ++p)
      for (int l = 0; l < 32; ++l) {
        const int r = s * rps + (l >> lg), k = p * G + (l & (G - 1));
        if (r < rows && H.rp[r] + k < H.rp[r + 1]) {
          sc[((size_t)sp[s] + p) * 32 + l] = H.col[H.rp[r] + k];
          sv[((size_t)sp[s] + p) * 32 + l] = H.val[H.rp[r] + k];
        }
      }
  const int gridS = OSQP_CUDA_SMS * SELL_MINB, nw = gridS * SELL_WARPS;
  std::vector<int> ws(nw + 1);
  for (int w = 0; w <= nw; ++w) {
    const long long target = (long long)tp * w / nw;
    ws[w] = (int)(std::lower_bound(sp.begin(), sp.end(), (int)target) - sp.begin());
    if (ws[w] > ns) ws[w] = ns;
  }
  ws[nw] = ns;
  SellView S{};
  std::vector<int> info(ns);
  for (int k = 0; k < ns; ++k) info[k] = ((k * rps) << 3) | lg;
  S.rows = rows; S.nslices = ns; S.split_col = split_col;
  S.slice_ptr = up(sp); S.slice_info = up(info); S.warp_slice = up(ws); S.col = up(sc, 256); S.val = up(sv, 256);
  const size_t smem = sell_smem_bytes();
  cudaFuncSetAttribute(spmv_sell_kernel<1, false, 0, EpiScale>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const float ts = timeit([&] { spmv_sell_kernel<1, false, 0, EpiScale><<<gridS, OSQP_CUDA_THREADS, smem>>>(S, g0, g1, e1, part); });
  cudaError_t err = cudaDeviceSynchronize();
  std::vector<double> h0(rows), h1(rows);
  cudaMemcpy(h0.data(), w0, rows * 8, cudaMemcpyDeviceToHost); cudaMemcpy(h1.data(), w1, rows * 8, cudaMemcpyDeviceToHost);
  double dmax = 0, amax = 0;
  for (int i = 0; i < rows; ++i) { dmax = std::max(dmax, std::fabs(h0[i] - h1[i])); amax = std::max(amax, std::fabs(h0[i])); }
  const double mb = (12.0 * nnz + 4.0 * rows + 8.0 * (nlo + nhi) + 16.0 * rows) / 1e6;
  printf("%-6s tma=%d pf=%d ch=%d nst=%d minb=%d lg=%d pad=%.3f | direct %7.2f us (%5.0f GB/s)  sell %7.2f us (%5.0f GB/s)  maxdiff %.2e (max|w| %.2e)  %s\n",
         name, SELL_TMA, SELL_PF, SELL_CH, SELL_NST, SELL_MINB, lg, (double)tp * 32 / nnz, td, mb / td * 1e3, ts, mb / ts * 1e3,
         dmax, amax, cudaGetErrorString(err));
}

int main() {
  const int n = 100000, m = 1000000;
  {
    Host A; A.rows = m; A.cols = n; A.rp.resize(m + 1);
    A.col.resize((size_t)m * 10); A.val.resize((size_t)m * 10);
    for (int r = 0; r <= m; ++r) A.rp[r] = r * 10;
    for (int r = 0; r < m; ++r) {
      int *c = &A.col[(size_t)r * 10];
      for (int k = 0; k < 10; ++k) c[k] = (int)(rnd() % n);
      std::sort(c, c + 10);
      for (int k = 0; k < 10; ++k) A.val[(size_t)r * 10 + k] = (double)(rnd() % 2001) / 1000.0 - 1.0;
    }
    run("A", A, 0, INT_MAX, n, 0);
  }
  {
    Host M; M.rows = n; M.cols = n + m; M.rp.resize(n + 1); M.rp[0] = 0;
    for (int r = 0; r < n; ++r) {
      const int la = 90 + (int)(rnd() % 21);
      std::vector<int> c;
      for (int k = 0; k < 7; ++k) c.push_back((int)(rnd() % n));
      std::sort(c.begin(), c.end());
      std::vector<int> c2;
      for (int k = 0; k < la; ++k) c2.push_back(n + (int)(rnd() % m));
      std::sort(c2.begin(), c2.end());
      c.insert(c.end(), c2.begin(), c2.end());
      for (int cc : c) { M.col.push_back(cc); M.val.push_back((double)(rnd() % 2001) / 1000.0 - 1.0); }
      M.rp[r + 1] = (int)M.col.size();
    }
    run("M", M, SELL_LGM, n, n, m);
  }
  return 0;
}
