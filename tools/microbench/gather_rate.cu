// Microbenchmark: peak rate of random 8-byte gathers from an L2-resident vector on B200,
// for different load flavours, table sizes and occupancies.  (Design input for the SpMV kernels.)
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
template <int MODE> __device__ __forceinline__ double ld(const double *p) {
  double v;
  if (MODE == 0) asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(v) : "l"(p));
  if (MODE == 1) asm volatile("ld.global.cg.f64 %0, [%1];" : "=d"(v) : "l"(p));
  if (MODE == 2) asm volatile("ld.global.ca.f64 %0, [%1];" : "=d"(v) : "l"(p));
  if (MODE == 3) asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
template <int MODE, int U>
__global__ void k_gather(const int *__restrict__ idx, const double *__restrict__ x, double *out, long long n) {
  double acc = 0.0;
  const long long stride = (long long)gridDim.x * blockDim.x * U;
  for (long long i = (long long)blockIdx.x * blockDim.x * U + threadIdx.x; i < n; i += stride) {
    int c[U]; double v[U];
#pragma unroll
    for (int k = 0; k < U; ++k) c[k] = (i + k * blockDim.x < n) ? idx[i + k * blockDim.x] : 0;
#pragma unroll
    for (int k = 0; k < U; ++k) v[k] = ld<MODE>(x + c[k]);
#pragma unroll
    for (int k = 0; k < U; ++k) acc += v[k];
  }
  if (acc == 1.2345e300) out[0] = acc;
}
template <int U>
__global__ void k_gather_smem(const int *__restrict__ idx, const double *__restrict__ x, double *out, long long n, int tab) {
  extern __shared__ double sx[];
  for (int i = threadIdx.x; i < tab; i += blockDim.x) sx[i] = x[i];
  __syncthreads();
  double acc = 0.0;
  const long long stride = (long long)gridDim.x * blockDim.x * U;
  for (long long i = (long long)blockIdx.x * blockDim.x * U + threadIdx.x; i < n; i += stride) {
#pragma unroll
    for (int k = 0; k < U; ++k) { long long j = i + k * blockDim.x; if (j < n) acc += sx[idx[j] % tab]; }
  }
  if (acc == 1.2345e300) out[0] = acc;
}
template <class F> float timeit(F f) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); f();
  cudaEventRecord(a);
  for (int r = 0; r < 5; ++r) f();
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms / 5;
}
int main() {
  const long long n = 40000000;
  int *idx; double *x, *out;
  cudaMalloc(&idx, n * 4); cudaMalloc(&out, 8);
  for (int tab : {100000, 1000000, 10000000}) {
    std::vector<int> h(n);
    unsigned long long s = 88172645463325252ull;
    for (long long i = 0; i < n; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (int)(s % tab); }
    cudaMemcpy(idx, h.data(), n * 4, cudaMemcpyHostToDevice);
    cudaMalloc(&x, (size_t)tab * 8); cudaMemset(x, 0, (size_t)tab * 8);
    for (int bps : {2, 4, 8}) {
      const int grid = 148 * bps, blk = 256;
#define RUN(MODE, U, name) { float ms = timeit([&] { k_gather<MODE, U><<<grid, blk>>>(idx, x, out, n); }); \
      printf("tab=%8d ctas/sm=%d %-22s U=%d : %7.3f ms  %6.1f Ggather/s  %.3f gathers/clk/SM (1.93GHz)\n", tab, bps, name, U, ms, n / ms / 1e6, n / (ms * 1e-3) / 148 / 1.93e9); }
      RUN(0, 8, "ld.global.nc") RUN(1, 8, "ld.global.cg") RUN(2, 8, "ld.global.ca") RUN(3, 8, "nc.L1::no_allocate") RUN(0, 16, "ld.global.nc") RUN(1, 16, "ld.global.cg")
    }
    if (tab == 100000) {
      const int stab = 25000;
      cudaFuncSetAttribute(k_gather_smem<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, stab * 8);
      float ms = timeit([&] { k_gather_smem<8><<<148, 1024, stab * 8>>>(idx, x, out, n, stab); });
      printf("smem table %d doubles, 1 CTA/SM x 1024 thr: %7.3f ms  %6.1f Ggather/s  %.3f gathers/clk/SM\n", stab, ms, n / ms / 1e6, n / (ms * 1e-3) / 148 / 1.93e9);
    }
    cudaFree(x);
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return 0;
}
