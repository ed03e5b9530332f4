// dev_common.cuh -- shared device/host helpers for libosqp_cuda_pcg.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cfloat>
#include <climits>
#include <vector>
#include "osqp_cuda_pcg.h"

#define OSQP_CUDA_THREADS 256          // threads per CTA for every kernel here
#define OSQP_CUDA_SMS     148          // B200: 2 dies x 74 SMs
#define OSQP_MAX_PARTIAL_BLOCKS 1184   // 148 x 8: upper bound on persistent grid size

#define CUDA_TRY(expr)                                                          \
  do {                                                                          \
    cudaError_t err__ = (expr);                                                 \
    if (err__ != cudaSuccess) {                                                 \
      fprintf(stderr, "osqp_cuda_pcg: CUDA error %s at %s:%d: %s\n",            \
              cudaGetErrorName(err__), __FILE__, __LINE__,                      \
              cudaGetErrorString(err__));                                       \
      return 1;                                                                 \
    }                                                                           \
  } while (0)

// Device allocations go through a small caching layer (dev_mats.cu): repeated
// osqp_setup / osqp_cleanup cycles (parametric re-solves, batches, bench e2e) ask for the
// same sizes again and again, and cudaMalloc / cudaFree of 100 MB buffers cost milliseconds.
int dev_cache_alloc(void **p, size_t bytes);
void dev_cache_free(void *p);

template <class T>
static inline int dev_alloc(T **p, size_t count) {
  *p = nullptr;
  size_t bytes = (count ? count : 1) * sizeof(T);
  void *q = nullptr;
  if (dev_cache_alloc(&q, bytes)) {
    fprintf(stderr, "osqp_cuda_pcg: device allocation of %zu bytes failed: %s\n", bytes,
            cudaGetErrorString(cudaGetLastError()));
    return 1;
  }
  *p = (T *)q;
  return 0;
}
template <class T>
static inline void dev_free(T *&p) {
  if (p) dev_cache_free((void *)p);
  p = nullptr;
}

// -------------------------------------------------------------- reductions
// All reductions are fixed-shape trees: the same inputs always give the same
// bits (no floating-point atomics anywhere on the solve path).

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum / max over OSQP_CUDA_THREADS threads; result valid in all threads.
// `scratch` is >= 32 doubles of shared memory; two __syncthreads inside.
template <bool IS_MAX>
__device__ __forceinline__ double block_reduce(double v, double *scratch) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  v = IS_MAX ? warp_max(v) : warp_sum(v);
  __syncthreads();                 // protect scratch from a previous use
  if (lane == 0) scratch[wid] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  double r = (lane < nw) ? scratch[lane] : (IS_MAX ? -DBL_MAX : 0.0);
  r = IS_MAX ? warp_max(r) : warp_sum(r);
  return r;
}

// Sum `count` per-block partials (stride `stride` doubles apart) redundantly in
// every block: each block gets the identical total without a host round trip.
__device__ __forceinline__ double reduce_partials_sum(const double *partials, int count,
                                                      int stride, double *scratch) {
  double v = 0.0;
  for (int i = threadIdx.x; i < count; i += blockDim.x) v += partials[(size_t)i * stride];
  return block_reduce<false>(v, scratch);
}

__device__ __forceinline__ double reduce_partials_max(const double *partials, int count,
                                                      int stride, double *scratch) {
  double v = -DBL_MAX;
  for (int i = threadIdx.x; i < count; i += blockDim.x) v = fmax(v, partials[(size_t)i * stride]);
  return block_reduce<true>(v, scratch);
}

// order-independent max on non-negative doubles (bit pattern is monotone)
__device__ __forceinline__ void atomic_max_nonneg(double *addr, double v) {
  atomicMax((unsigned long long *)addr, (unsigned long long)__double_as_longlong(v));
}

static inline int grid_for(long long work_items, int per_block = OSQP_CUDA_THREADS) {
  long long b = (work_items + per_block - 1) / per_block;
  if (b < 1) b = 1;
  if (b > 65535LL * 16) b = 65535LL * 16;
  return (int)b;
}
