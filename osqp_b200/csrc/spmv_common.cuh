// spmv_common.cuh -- what every SpMV kernel of the product shares: the device CSR descriptor with its
// mirrors (sliced ELL: spmv_sell.cuh, column-blocked: spmv_tile.cuh), the gather bases, the row-epilogue
// concept, TMA / mbarrier helpers.
//
// The SpMVs replace the CSC scatter/gather loops of reference src/lin_alg.c:241-322 (mat_vec /
// mat_tpose_vec) and, through the row epilogues, the vector loops of src/auxil.c:161-225 that consume
// their results.  (Round 1's warp-specialised TMA "stream" kernel lived here; it lost to the direct and
// sliced-ELL kernels on every workload and is kept for the record under tools/microbench/spmv_stream.cuh.)
#pragma once
#include "dev_common.cuh"

#define SPMV_TILE      1024   // nonzeros staged per tile
#define SPMV_CHUNK_NNZ 896    // chunk quantum; a chunk spans < CHUNK + longest row, so rows of
                              // up to TILE - CHUNK = 128 entries keep a chunk inside ONE tile
#define SPMV_TILE_PAD  (SPMV_TILE + 8)
#define SPMV_RP_MAX    (SPMV_TILE + 8)   // row pointers of a chunk staged in shared memory
#define SPMV_PER_THREAD (SPMV_TILE / OSQP_CUDA_THREADS)

// Kernel selection (experiment knob, read once): OSQP_B200_SPMV=direct keeps the SpMVs on the CSR arrays
// (spmv_direct.cuh) instead of the mirrors.
#include <cstdlib>
#include <cstring>
static inline int spmv_kind() {   // 1 direct CSR, 2 mirrors (default)
  static int sel = -1;
  if (sel < 0) {
    const char *e = getenv("OSQP_B200_SPMV");
    sel = (e && !strcmp(e, "direct")) ? 1 : 2;
  }
  return sel;
}

// sliced-ELL mirror of a CSR matrix (layout and kernel: spmv_sell.cuh; built by dev_mats.cu)
struct SellDev {
  bool ok = false;          // mirror exists and is worth using
  bool dirty = true;        // CSR values changed since the last sell_refresh
  int nslices = 0, total_passes = 0, grid = 0;
  int *slice_ptr = nullptr, *slice_info = nullptr, *warp_slice = nullptr, *col = nullptr;
  double *val = nullptr;
};

// column-blocked shared-memory-gather mirror (layouts and kernels: spmv_tile.cuh; built by dev_mats.cu)
struct TileDev {
  bool ok = false;          // mirror exists and is worth using
  bool dirty = true;        // CSR values changed since the last tile_refresh
  bool rows_mode = false;   // true: CTA owns rows across all blocks (A); false: partial sums per block (A' part of M)
  int n_lo = 0, cb_lo = 0, nb_lo = 0;   // first column range (P of M): nb_lo blocks of cb_lo; then blocks of cb
  int rows = 0, rows_pad = 0, cols = 0, cb = 0, nblocks = 0, nwin = 0, win = 0, ncta = 0, nitems = 0;
  int racc = 0;             // ROWS mode: slices of the largest CTA * 32 = shared-memory accumulators
  int nsb = 0;              // slices per column block
  int total_passes = 0;
  int *slice_ptr = nullptr, *item_ptr = nullptr, *item_blk = nullptr, *item_ws = nullptr;
  int *cta_row = nullptr, *cta_soff = nullptr;
  int2 *ne = nullptr;                         // slices with entries, in order: {first pass, slice number} (the kernels walk these)
  int *wb_ptr = nullptr, *wb_blk = nullptr;   // PARTIAL: per window, the column blocks in which it has entries
  int *src = nullptr;                 // [passes * 32]: CSR slot of the entry, -1 = padding
  unsigned short *perm = nullptr;     // ROWS: sorted position -> row in window;  PARTIAL: row -> sorted position
  unsigned short *col = nullptr;
  double *val = nullptr;
  double *partial = nullptr;          // PARTIAL: [nblocks][rows_pad] partial row sums
};

struct CsrDev {
  int rows = 0, cols = 0, nnz = 0;
  int *rowptr = nullptr;     // rows+1
  int *col = nullptr;        // nnz, padded to a multiple of 4 (+4), pad = 0
  double *val = nullptr;     // same padding, pad = 0.0
  int *rowidx = nullptr;     // nnz: row of each entry (setup / scaling kernels)
  int *src = nullptr;        // nnz: slot in the caller's CSC value array
  int *split = nullptr;      // rows: first A' entry of row j in [P | A'] (null for A)
  int nchunks = 0;
  int *chunk_row = nullptr;  // nchunks+1
  int4 *chunk_desc = nullptr; // nchunks: {first row, end row, rowptr[first], rowptr[end]}
  int group = 1;             // lanes per row in the reduction (power of two)
  int split_col = INT_MAX;   // columns >= split_col gather from the second base
  mutable SellDev sell;      // the layout the large-problem SpMV launches stream
  mutable TileDev tile;      // column-blocked mirror for the hot SpMVs of large random-sparsity problems
};

struct CsrView {
  int rows, nnz, nchunks, group, split_col;
  const int *rowptr, *col, *chunk_row, *split, *rowidx;
  const double *val;
  const int4 *chunk_desc;
};
static inline CsrView view_of(const CsrDev &A) {
  CsrView v;
  v.rows = A.rows; v.nnz = A.nnz; v.nchunks = A.nchunks; v.group = A.group;
  v.split_col = A.split_col; v.rowptr = A.rowptr; v.col = A.col;
  v.chunk_row = A.chunk_row; v.split = A.split; v.rowidx = A.rowidx; v.val = A.val;
  v.chunk_desc = A.chunk_desc;
  return v;
}

// Gather bases: column c < split_col reads lo[c], else hi[c - split_col].
struct Gather {
  const double *lo;
  const double *hi;
};

// MODE 0: product = val * gathered
// MODE 1: Jacobi diagonal of P + A' diag(rho) A on [P | A']:
//         P part contributes val on the diagonal, A' part val^2 * rho[c - n]
template <int MODE>
__device__ __forceinline__ double spmv_product(double v, int c, int row_of_entry,
                                               const Gather &g, int split_col) {
  if (MODE == 0) {
    const double xv = (c < split_col) ? __ldg(g.lo + c) : __ldg(g.hi + (c - split_col));
    return v * xv;
  } else {
    if (c < split_col) return (c == row_of_entry) ? v : 0.0;
    return v * v * __ldg(g.hi + (c - split_col));
  }
}

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gmem_src, unsigned bytes,
                                         unsigned long long *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
               ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

// Epilogue concept:
//   static constexpr int NSUM, NMAX;          per-thread accumulators reduced per CTA
//   __device__ bool begin(double *scratch);   CTA-uniform prologue; false = skip kernel
//   struct Pre; __device__ Pre load(int r) const;   per-row inputs, loaded BEFORE the row
//                                             reduction so their latency overlaps it
//   __device__ void row(int r, const double *s, const Pre &pre, double *asum, double *amax) const;
// where s[v * NS + part] is the row sum of gathered vector v (part 1 = A' part
// when SPLIT).  Partials land in partials[blockIdx.x * (NSUM + NMAX) + k].
// Experiment switch (tools/spmv_phases.py): bit 0 = skip the gathers, bit 1 = skip the row
// reduction.  Zero in production; a uniform branch on a __device__ variable.
static __device__ int g_spmv_dbg = 0;
// non-blocking gather: the load is issued here, the value is consumed later
__device__ __forceinline__ double ldg_issue(const double *p) {
  double v;
  asm volatile("ld.global.nc.f64 %0, [%1];\n" : "=d"(v) : "l"(p) : "memory");
  return v;
}
#define EPI_NO_PREFETCH \
  struct Pre {};        \
  __device__ __forceinline__ Pre load(int) const { return Pre(); }

__device__ __forceinline__ void mbar_arrive(unsigned long long *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}

// Persistent grid: a multiple of the SM count, capped by the chunk count.
static inline int spmv_grid(const CsrDev &A, int ctas_per_sm) {
  if (A.sell.ok) return A.sell.grid;   // the mirror's warp -> slice table is built for this grid
  const char *e = getenv("OSQP_B200_CPS");   // experiment knob: resident CTAs per SM
  if (e && atoi(e) > 0) ctas_per_sm = atoi(e);
  int g = OSQP_CUDA_SMS * ctas_per_sm;
  if (g > OSQP_MAX_PARTIAL_BLOCKS) g = OSQP_MAX_PARTIAL_BLOCKS;
  if (g > A.nchunks) g = A.nchunks;
  if (g < 1) g = 1;
  return g;
}

