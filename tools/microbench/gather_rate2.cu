// Microbenchmark (round 2): how fast can one SM gather / scatter random fp64 elements when the
// table is NOT behind L1TEX?  Replaces the shared-memory arm of gather_rate.cu, which measured an
// integer modulo (sx[idx % tab]), a serial add chain and the table fill instead of shared memory.
//
// Arms (all indices are pre-reduced on the host, tables are filled before the timed loop, U
// independent loads are issued before any add):
//   smem_pure   : ld.shared.f64 on 32-bit addresses, every loaded index reused R times with a
//                 power-of-two table -> the shared-memory gather rate with no index stream
//   smem_spmv   : the SpMV inner loop: stream (col, val) from HBM, gather x[col] from shared
//                 memory, fma.  col as int32 (12 B/nnz) or uint16 (10 B/nnz)
//   l2_spmv     : the same loop gathering from global memory (L1TEX path) -- the baseline
//   dsmem_spmv  : the table spread over the shared memory of a CL-CTA cluster (CL = 2, 4, 8),
//                 gathers through mapa + ld.shared::cluster.f64
//   smem_red / dsmem_red / l2_red : random fp64 scatter-adds (red.shared / red.shared::cluster /
//                 red.global) -- what a transposed product without an explicit A' would need
// Prints gathers/clk/SM at the measured event time and 1.965 GHz.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>

#define CHECK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)
static const double GHZ = 1.965;

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ double lds64(unsigned a) { double v; asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v; }
__device__ __forceinline__ unsigned mapa(unsigned a, unsigned rank) { unsigned r; asm("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank)); return r; }
__device__ __forceinline__ double ldsc64(unsigned a) { double v; asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v; }
__device__ __forceinline__ void cluster_sync() { asm volatile("barrier.cluster.arrive.aligned;\nbarrier.cluster.wait.aligned;\n" ::: "memory"); }
__device__ __forceinline__ unsigned cluster_rank() { unsigned r; asm("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }

template <int U, int R>
__global__ void __launch_bounds__(1024, 1)
k_smem_pure(const int *__restrict__ idx, const double *__restrict__ x, double *out, long long n, int tab) {
  extern __shared__ double sx[];
  for (int i = threadIdx.x; i < tab; i += blockDim.x) sx[i] = x[i];
  __syncthreads();
  const unsigned base = smem_u32(sx), mask = (unsigned)tab - 1u;
  double acc[U];
#pragma unroll
  for (int k = 0; k < U; ++k) acc[k] = 0.0;
  const long long stride = (long long)gridDim.x * blockDim.x * U;
  for (long long i = (long long)blockIdx.x * blockDim.x * U + threadIdx.x; i + (U - 1) * blockDim.x < n; i += stride) {
    unsigned c[U];
#pragma unroll
    for (int k = 0; k < U; ++k) c[k] = (unsigned)idx[i + k * blockDim.x];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      double t[U];
#pragma unroll
      for (int k = 0; k < U; ++k) t[k] = lds64(base + 8u * ((c[k] + 7919u * r) & mask));
#pragma unroll
      for (int k = 0; k < U; ++k) acc[k] += t[k];
    }
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < U; ++k) s += acc[k];
  if (s == 1.2345e300) out[0] = s;
}

// SRC 0: table in shared memory (tab doubles per CTA); SRC 1: table in global memory
template <int U, class IDX, int SRC>
__global__ void __launch_bounds__(1024, 1)
k_spmv_like(const IDX *__restrict__ col, const double *__restrict__ val, const double *__restrict__ x,
            double *out, long long n, int tab) {
  extern __shared__ double sx[];
  if (SRC == 0) {
    for (int i = threadIdx.x; i < tab; i += blockDim.x) sx[i] = x[i];
    __syncthreads();
  }
  const unsigned base = smem_u32(sx);
  double acc[U];
#pragma unroll
  for (int k = 0; k < U; ++k) acc[k] = 0.0;
  const long long stride = (long long)gridDim.x * blockDim.x * U;
  for (long long i = (long long)blockIdx.x * blockDim.x * U + threadIdx.x; i + (U - 1) * blockDim.x < n; i += stride) {
    unsigned c[U];
    double v[U], t[U];
#pragma unroll
    for (int k = 0; k < U; ++k) c[k] = (unsigned)__ldcs(col + i + k * blockDim.x);
#pragma unroll
    for (int k = 0; k < U; ++k) v[k] = __ldcs(val + i + k * blockDim.x);
#pragma unroll
    for (int k = 0; k < U; ++k) t[k] = (SRC == 0) ? lds64(base + 8u * c[k]) : __ldg(x + c[k]);
#pragma unroll
    for (int k = 0; k < U; ++k) acc[k] = fma(v[k], t[k], acc[k]);
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < U; ++k) s += acc[k];
  if (s == 1.2345e300) out[0] = s;
}

// table of CL * tabl doubles spread over the cluster; col packed as (rank << 20) | offset
template <int U>
__global__ void __launch_bounds__(1024, 1)
k_dsmem_spmv(const int *__restrict__ col, const double *__restrict__ val, const double *__restrict__ x,
             double *out, long long n, int tabl) {
  extern __shared__ double sx[];
  const unsigned rank = cluster_rank();
  for (int i = threadIdx.x; i < tabl; i += blockDim.x) sx[i] = x[(size_t)rank * tabl + i];
  cluster_sync();
  const unsigned base = smem_u32(sx);
  double acc[U];
#pragma unroll
  for (int k = 0; k < U; ++k) acc[k] = 0.0;
  const long long stride = (long long)gridDim.x * blockDim.x * U;
  for (long long i = (long long)blockIdx.x * blockDim.x * U + threadIdx.x; i + (U - 1) * blockDim.x < n; i += stride) {
    unsigned c[U];
    double v[U], t[U];
#pragma unroll
    for (int k = 0; k < U; ++k) c[k] = (unsigned)__ldcs(col + i + k * blockDim.x);
#pragma unroll
    for (int k = 0; k < U; ++k) v[k] = __ldcs(val + i + k * blockDim.x);
#pragma unroll
    for (int k = 0; k < U; ++k) t[k] = ldsc64(mapa(base + 8u * (c[k] & 0xfffffu), c[k] >> 20));
#pragma unroll
    for (int k = 0; k < U; ++k) acc[k] = fma(v[k], t[k], acc[k]);
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < U; ++k) s += acc[k];
  if (s == 1.2345e300) out[0] = s;
  cluster_sync();   // nobody leaves while a peer may still read its shared memory
}

// scatter-adds.  DST 0: local shared memory, 1: cluster shared memory (packed col), 2: global
template <int U, int DST>
__global__ void __launch_bounds__(1024, 1)
k_red(const int *__restrict__ col, const double *__restrict__ val, double *table, double *out, long long n, int tabl) {
  extern __shared__ double sx[];
  if (DST != 2) for (int i = threadIdx.x; i < tabl; i += blockDim.x) sx[i] = 0.0;
  if (DST == 1) cluster_sync(); else __syncthreads();
  const unsigned base = smem_u32(sx);
  const long long stride = (long long)gridDim.x * blockDim.x * U;
  for (long long i = (long long)blockIdx.x * blockDim.x * U + threadIdx.x; i + (U - 1) * blockDim.x < n; i += stride) {
    unsigned c[U];
    double v[U];
#pragma unroll
    for (int k = 0; k < U; ++k) c[k] = (unsigned)__ldcs(col + i + k * blockDim.x);
#pragma unroll
    for (int k = 0; k < U; ++k) v[k] = __ldcs(val + i + k * blockDim.x);
#pragma unroll
    for (int k = 0; k < U; ++k) {
      if (DST == 0) asm volatile("red.shared.add.f64 [%0], %1;" ::"r"(base + 8u * c[k]), "d"(v[k]) : "memory");
      if (DST == 1) asm volatile("red.shared::cluster.add.f64 [%0], %1;" ::"r"(mapa(base + 8u * (c[k] & 0xfffffu), c[k] >> 20)), "d"(v[k]) : "memory");
      if (DST == 2) asm volatile("red.global.add.f64 [%0], %1;" ::"l"(table + c[k]), "d"(v[k]) : "memory");
    }
  }
  if (DST == 1) cluster_sync(); else __syncthreads();
  if (DST != 2 && sx[threadIdx.x % tabl] == 1.2345e300) out[0] = 1.0;
}

template <class F> float timeit(F f, int reps = 5) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); f();
  CHECK(cudaDeviceSynchronize());
  cudaEventRecord(a);
  for (int r = 0; r < reps; ++r) f();
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  CHECK(cudaGetLastError());
  return ms / reps;
}
static unsigned long long rs = 88172645463325252ull;
static inline unsigned long long rnd() { rs ^= rs << 13; rs ^= rs >> 7; rs ^= rs << 17; return rs; }

static void report(const char *name, long long gathers, float ms, int sms, double bytes_per = 0.0) {
  printf("%-44s %8.3f ms  %7.1f G/s  %6.3f per clk per SM (%d SMs, %.3f GHz)", name, ms, gathers / ms / 1e6,
         gathers / (ms * 1e-3) / sms / (GHZ * 1e9), sms, GHZ);
  if (bytes_per > 0) printf("  stream %.0f GB/s", gathers * bytes_per / ms / 1e6);
  printf("\n");
  fflush(stdout);
}

template <int CL, int U>
static void run_cluster(const char *what, bool red, const int *dcol, const double *dval, const double *dx, double *dtab,
                        double *dout, long long n, int tabl) {
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cfg.blockDim = dim3(1024); cfg.dynamicSmemBytes = (size_t)tabl * 8;
  int nclusters = 0;
  cfg.gridDim = dim3(CL * 64);
  if (red) {
    auto *kern = k_red<U, 1>;
    CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tabl * 8));
    if (CL > 8) CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    CHECK(cudaOccupancyMaxActiveClusters(&nclusters, kern, &cfg));
    cfg.gridDim = dim3(nclusters * CL);
    float ms = timeit([&] { CHECK(cudaLaunchKernelEx(&cfg, kern, dcol, dval, dtab, dout, n, tabl)); });
    char nm[128]; snprintf(nm, sizeof nm, "%s CL=%d tab/CTA=%d clusters=%d U=%d", what, CL, tabl, nclusters, U);
    report(nm, n, ms, nclusters * CL, 12.0);
  } else {
    auto *kern = k_dsmem_spmv<U>;
    CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tabl * 8));
    if (CL > 8) CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    CHECK(cudaOccupancyMaxActiveClusters(&nclusters, kern, &cfg));
    cfg.gridDim = dim3(nclusters * CL);
    float ms = timeit([&] { CHECK(cudaLaunchKernelEx(&cfg, kern, dcol, dval, dx, dout, n, tabl)); });
    char nm[128]; snprintf(nm, sizeof nm, "%s CL=%d tab/CTA=%d clusters=%d U=%d", what, CL, tabl, nclusters, U);
    report(nm, n, ms, nclusters * CL, 12.0);
  }
}

int main() {
  const long long n = 40ll * 1024 * 1024;   // gathers per launch (multiple of 1024 * U * 148? no: tail ignored)
  int *dcol; unsigned short *dcol16; double *dval, *dx, *dout, *dtab;
  CHECK(cudaMalloc(&dcol, n * 4)); CHECK(cudaMalloc(&dcol16, n * 2)); CHECK(cudaMalloc(&dval, n * 8));
  CHECK(cudaMalloc(&dx, 8 << 20)); CHECK(cudaMalloc(&dout, 64)); CHECK(cudaMalloc(&dtab, 8 << 20));
  CHECK(cudaMemset(dx, 0, 8 << 20)); CHECK(cudaMemset(dval, 0, n * 8)); CHECK(cudaMemset(dtab, 0, 8 << 20));
  std::vector<int> h(n); std::vector<unsigned short> h16(n);
  auto fill = [&](int tab, int tabl /* per-CTA part for packed indices, 0 = plain */) {
    for (long long i = 0; i < n; ++i) {
      const int c = (int)(rnd() % (unsigned)tab);
      h[i] = tabl ? (((c / tabl) << 20) | (c % tabl)) : c;
      h16[i] = (unsigned short)(c & 0xffff);
    }
    CHECK(cudaMemcpy(dcol, h.data(), n * 4, cudaMemcpyHostToDevice));
    CHECK(cudaMemcpy(dcol16, h16.data(), n * 2, cudaMemcpyHostToDevice));
  };

  // ---- 1. pure shared-memory gather rate (power-of-two table, indices reused R = 8 times)
  {
    const int tab = 16384;
    fill(tab, 0);
    CHECK(cudaFuncSetAttribute(k_smem_pure<8, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, tab * 8));
    CHECK(cudaFuncSetAttribute(k_smem_pure<4, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, tab * 8));
    float ms = timeit([&] { k_smem_pure<8, 8><<<148, 1024, tab * 8>>>(dcol, dx, dout, n, tab); });
    report("smem_pure  tab=16384 U=8 R=8 1024thr", n * 8, ms, 148);
    ms = timeit([&] { k_smem_pure<4, 8><<<148, 1024, tab * 8>>>(dcol, dx, dout, n, tab); });
    report("smem_pure  tab=16384 U=4 R=8 1024thr", n * 8, ms, 148);
    ms = timeit([&] { k_smem_pure<8, 8><<<148, 512, tab * 8>>>(dcol, dx, dout, n, tab); });
    report("smem_pure  tab=16384 U=8 R=8 512thr", n * 8, ms, 148);
  }
  // ---- 2. SpMV-like loop, table in shared memory (25000 doubles = 200 KB) vs global memory
  {
    const int tab = 25000;
    fill(tab, 0);
#define SMEM_ATTR(K) CHECK(cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, tab * 8))
    SMEM_ATTR((k_spmv_like<4, int, 0>)); SMEM_ATTR((k_spmv_like<8, int, 0>)); SMEM_ATTR((k_spmv_like<4, unsigned short, 0>));
    SMEM_ATTR((k_spmv_like<8, unsigned short, 0>));
    float ms = timeit([&] { k_spmv_like<4, int, 0><<<148, 1024, tab * 8>>>(dcol, dval, dx, dout, n, tab); });
    report("smem_spmv  tab=25000 col=i32 U=4", n, ms, 148, 12.0);
    ms = timeit([&] { k_spmv_like<8, int, 0><<<148, 1024, tab * 8>>>(dcol, dval, dx, dout, n, tab); });
    report("smem_spmv  tab=25000 col=i32 U=8", n, ms, 148, 12.0);
    ms = timeit([&] { k_spmv_like<4, unsigned short, 0><<<148, 1024, tab * 8>>>(dcol16, dval, dx, dout, n, tab); });
    report("smem_spmv  tab=25000 col=u16 U=4", n, ms, 148, 10.0);
    ms = timeit([&] { k_spmv_like<8, unsigned short, 0><<<148, 1024, tab * 8>>>(dcol16, dval, dx, dout, n, tab); });
    report("smem_spmv  tab=25000 col=u16 U=8", n, ms, 148, 10.0);
    ms = timeit([&] { k_spmv_like<4, int, 0><<<148, 512, tab * 8>>>(dcol, dval, dx, dout, n, tab); });
    report("smem_spmv  tab=25000 col=i32 U=4 512thr", n, ms, 148, 12.0);
    // fill only (n = 0): cost of staging the 200 KB table per CTA
    ms = timeit([&] { k_spmv_like<4, int, 0><<<148, 1024, tab * 8>>>(dcol, dval, dx, dout, 0, tab); });
    printf("   (table fill only, 200 KB per CTA: %.2f us)\n", ms * 1e3);
    fill(100000, 0);
    ms = timeit([&] { k_spmv_like<4, int, 1><<<148, 1024, 0>>>(dcol, dval, dx, dout, n, 100000); });
    report("l2_spmv    tab=100000 col=i32 U=4 1x1024", n, ms, 148, 12.0);
    ms = timeit([&] { k_spmv_like<4, int, 1><<<296, 1024, 0>>>(dcol, dval, dx, dout, n, 100000); });
    report("l2_spmv    tab=100000 col=i32 U=4 2x1024", n, ms, 148, 12.0);
    ms = timeit([&] { k_spmv_like<8, int, 1><<<296, 1024, 0>>>(dcol, dval, dx, dout, n, 100000); });
    report("l2_spmv    tab=100000 col=i32 U=8 2x1024", n, ms, 148, 12.0);
  }
  // ---- 3. table spread over a cluster's shared memory (whole 800 KB vector for CL >= 4)
  {
    fill(50000, 25000);
    run_cluster<2, 4>("dsmem_spmv tab=50000 ", false, dcol, dval, dx, dtab, dout, n, 25000);
    fill(100000, 25000);
    run_cluster<4, 4>("dsmem_spmv tab=100000", false, dcol, dval, dx, dtab, dout, n, 25000);
    run_cluster<4, 8>("dsmem_spmv tab=100000", false, dcol, dval, dx, dtab, dout, n, 25000);
    fill(100000, 12500);
    run_cluster<8, 4>("dsmem_spmv tab=100000", false, dcol, dval, dx, dtab, dout, n, 12500);
    run_cluster<8, 8>("dsmem_spmv tab=100000", false, dcol, dval, dx, dtab, dout, n, 12500);
    fill(100000, 6250);
    run_cluster<16, 4>("dsmem_spmv tab=100000", false, dcol, dval, dx, dtab, dout, n, 6250);
  }
  // ---- 4. scatter-adds
  {
    const int tab = 25000;
    fill(tab, 0);
    CHECK(cudaFuncSetAttribute(k_red<4, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, tab * 8));
    float ms = timeit([&] { k_red<4, 0><<<148, 1024, tab * 8>>>(dcol, dval, dtab, dout, n, tab); });
    report("smem_red   tab=25000 U=4", n, ms, 148, 12.0);
    fill(100000, 0);
    ms = timeit([&] { k_red<4, 2><<<296, 1024, 0>>>(dcol, dval, dtab, dout, n, 1); });
    report("l2_red     tab=100000 U=4 2x1024", n, ms, 148, 12.0);
    fill(1000000, 0);
    ms = timeit([&] { k_red<4, 2><<<296, 1024, 0>>>(dcol, dval, dtab, dout, n, 1); });
    report("l2_red     tab=1000000 U=4 2x1024", n, ms, 148, 12.0);
    fill(100000, 25000);
    run_cluster<4, 4>("dsmem_red  tab=100000", true, dcol, dval, dx, dtab, dout, n, 25000);
    fill(100000, 12500);
    run_cluster<8, 4>("dsmem_red  tab=100000", true, dcol, dval, dx, dtab, dout, n, 12500);
  }
  CHECK(cudaDeviceSynchronize());
  printf("done\n");
  return 0;
}
