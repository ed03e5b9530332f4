// dist.cuh -- row-partitioned mode: one process per GPU, GPU g owns a slab of
// the rows of A (and l, u, z, y, rho, E); x, q, D, the CG n-vectors and all scalars are
// replicated; P's nonzeros are spread over the ranks (any split whose symmetric completions sum to
// P works: osqp_b200/dist.py hands rank g the g-th slab of columns of triu(P)).  Every SpMV over
// M_g = [P_g | A_g'] yields a partial n-vector that is joined by ncclAllReduce(sum) BEFORE
// its row epilogue runs, so the fused epilogues are reused unchanged (k_apply_epi).
// NCCL is loaded with dlopen at run time: the library has no link-time NCCL dependency
// and single-GPU users never load it.
#pragma once
#include <nccl.h>
#include <dlfcn.h>
#include <cstdio>
#include <cstring>

struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

struct DistCtx {
  ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
  bool loopback = false;      // world of one WITH the partitioned code path (collectives are identities): tests
  long long allreduces = 0;   // collectives enqueued (statistics)
};

static NcclApi g_nccl;
static DistCtx g_dist;        // process-wide: one process drives one GPU

static int nccl_load() {
  if (g_nccl.handle) return 0;
  const char *names[] = {getenv("OSQP_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  void *h = nullptr;
  for (const char *nm : names) {
    if (nm && *nm && (h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL))) break;
  }
  if (!h) {
    fprintf(stderr, "osqp_cuda_pcg: cannot load NCCL (%s)\n", dlerror());
    return 1;
  }
  NcclApi a;
  a.handle = h;
  a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(h, "ncclGetUniqueId");
  a.CommInitRank = (decltype(a.CommInitRank))dlsym(h, "ncclCommInitRank");
  a.CommDestroy = (decltype(a.CommDestroy))dlsym(h, "ncclCommDestroy");
  a.AllReduce = (decltype(a.AllReduce))dlsym(h, "ncclAllReduce");
  a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
  if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllReduce || !a.GetErrorString) {
    fprintf(stderr, "osqp_cuda_pcg: the NCCL library lacks a required symbol\n");
    dlclose(h);
    return 1;
  }
  g_nccl = a;
  return 0;
}

// sum or max of `count` doubles (or ints) across the ranks, in place, on stream st
static inline int dist_allreduce(DistCtx *d, void *buf, size_t count, bool is_max, bool is_int,
                                 cudaStream_t st) {
  if (!d || count == 0) return 0;
  if (d->world <= 1) { d->allreduces += d->loopback ? 1 : 0; return 0; }
  const ncclResult_t r = g_nccl.AllReduce(buf, buf, count, is_int ? ncclInt32 : ncclFloat64,
                                          is_max ? ncclMax : ncclSum, d->comm, st);
  d->allreduces += 1;
  if (r != ncclSuccess) {
    fprintf(stderr, "osqp_cuda_pcg: ncclAllReduce failed: %s\n", g_nccl.GetErrorString(r));
    return 1;
  }
  return 0;
}

// ---- "store the row sums" epilogue + "apply the real epilogue to reduced sums" kernel
template <int NACC>
struct EpiStoreSums {
  static constexpr int NSUM = 0, NMAX = 0;
  double *out;             // rows x NACC
  const int *skip;         // device flag: nonzero = this launch is a no-op (CG already converged)
  __device__ bool begin(double *) { return !(skip && *skip); }
  EPI_NO_PREFETCH
  __device__ void row(int r, const double *s, const Pre &, double *, double *) const {
#pragma unroll
    for (int k = 0; k < NACC; ++k) out[(size_t)r * NACC + k] = s[k];
  }
};

template <int NACC, class Epi>
__global__ void __launch_bounds__(OSQP_CUDA_THREADS)
k_apply_epi(Epi epi_in, const double *sums, int rows, double *partials) {
  __shared__ double scratch[32];
  Epi epi = epi_in;
  if (!epi.begin(scratch)) return;
  double asum[Epi::NSUM > 0 ? Epi::NSUM : 1];
  double amax[Epi::NMAX > 0 ? Epi::NMAX : 1];
#pragma unroll
  for (int k = 0; k < Epi::NSUM; ++k) asum[k] = 0.0;
#pragma unroll
  for (int k = 0; k < Epi::NMAX; ++k) amax[k] = -DBL_MAX;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += gridDim.x * blockDim.x) {
    const typename Epi::Pre pre = epi.load(r);
    double s[NACC];
#pragma unroll
    for (int k = 0; k < NACC; ++k) s[k] = sums[(size_t)r * NACC + k];
    epi.row(r, s, pre, asum, amax);
  }
  if (Epi::NSUM + Epi::NMAX > 0) {
    double *out = partials + (size_t)blockIdx.x * (Epi::NSUM + Epi::NMAX);
#pragma unroll
    for (int k = 0; k < Epi::NSUM; ++k) {
      const double t = block_reduce<false>(asum[k], scratch);
      if (threadIdx.x == 0) out[k] = t;
    }
#pragma unroll
    for (int k = 0; k < Epi::NMAX; ++k) {
      const double t = block_reduce<true>(amax[k], scratch);
      if (threadIdx.x == 0) out[Epi::NSUM + k] = t;
    }
  }
}

// per-CTA partials [nparts x (nsum + nmax)] -> one row [nsum sums | nmax maxes]  (one CTA)
__global__ void __launch_bounds__(OSQP_CUDA_THREADS)
k_compact_partials(const double *part, int nparts, int nsum, int nmax, double *out) {
  __shared__ double scratch[32];
  const int stride = nsum + nmax;
  for (int k = 0; k < nsum; ++k) {
    const double t = reduce_partials_sum(part + k, nparts, stride, scratch);
    if (threadIdx.x == 0) out[k] = t;
  }
  for (int k = 0; k < nmax; ++k) {
    const double t = reduce_partials_max(part + nsum + k, nparts, stride, scratch);
    if (threadIdx.x == 0) out[nsum + k] = t;
  }
}
