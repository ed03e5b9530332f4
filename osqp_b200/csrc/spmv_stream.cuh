// spmv_stream.cuh -- row-wise CSR SpMV for B200, "stream through shared memory".
//
// Replaces the CSC scatter/gather loops of reference src/lin_alg.c:241-322
// (mat_vec / mat_tpose_vec) and, through the row epilogues, the vector loops of
// src/auxil.c:161-225 that consume their results.
//
// Design (HBM-bound, no tensor cores):
//  * the nonzeros of a CSR matrix are cut into CHUNKS of whole rows: chunk k owns
//    the rows whose first nonzero lies in [k*CHUNK_NNZ, (k+1)*CHUNK_NNZ).  A
//    persistent grid (a multiple of the 148 SMs) walks the chunks round-robin,
//    so the work split depends on the matrix only -> bit-reproducible sums;
//  * a CTA streams a tile of val[]/col[] with 128-bit coalesced loads (int4 of
//    column indices + 2 x double2 of values per thread per step, independent of
//    row boundaries), multiplies by the gathered vector entries (L1/L2 hits) and
//    parks the products in shared memory;
//  * groups of G lanes (G = 1..32 chosen from the mean row length: "vector per
//    row" / "warp per row") then reduce each row segment out of shared memory in
//    a fixed order and run a fused row epilogue (rho scaling, ADMM updates,
//    residual norms, dot-product partials ...);
//  * rows longer than a tile are carried across tiles inside the CTA.
#pragma once
#include "dev_common.cuh"

#define SPMV_TILE      3072   // products staged per tile (per gathered vector)
#define SPMV_CHUNK_NNZ 2048   // chunk quantum; a chunk spans < CHUNK + longest row
#define SPMV_TILE_PAD  (SPMV_TILE + 4)

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
  int group = 1;             // lanes per row in the reduction (power of two)
  int split_col = INT_MAX;   // columns >= split_col gather from the second base
};

struct CsrView {
  int rows, nnz, nchunks, group, split_col;
  const int *rowptr, *col, *chunk_row, *split, *rowidx;
  const double *val;
};
static inline CsrView view_of(const CsrDev &A) {
  CsrView v;
  v.rows = A.rows; v.nnz = A.nnz; v.nchunks = A.nchunks; v.group = A.group;
  v.split_col = A.split_col; v.rowptr = A.rowptr; v.col = A.col;
  v.chunk_row = A.chunk_row; v.split = A.split; v.rowidx = A.rowidx; v.val = A.val;
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

// Epilogue concept:
//   static constexpr int NSUM, NMAX;          per-thread accumulators reduced per CTA
//   __device__ bool begin(double *scratch);   CTA-uniform prologue; false = skip kernel
//   __device__ void row(int r, const double *s, double *asum, double *amax) const;
// where s[v * NS + part] is the row sum of gathered vector v (part 1 = A' part
// when SPLIT).  Partials land in partials[blockIdx.x * (NSUM + NMAX) + k].
template <int NV, bool SPLIT, int MODE, class Epi>
__global__ void __launch_bounds__(OSQP_CUDA_THREADS)
spmv_stream_kernel(CsrView A, Gather g0, Gather g1, Epi epi_in, double *partials) {
  constexpr int NS = SPLIT ? 2 : 1;
  constexpr int NACC = NV * NS;
  extern __shared__ double smem[];
  double *prod = smem;                                  // NV * SPMV_TILE_PAD
  double *carry = smem + NV * SPMV_TILE_PAD;            // 2 * NACC
  double *scratch = carry + 2 * NACC;                   // 32

  Epi epi = epi_in;
  if (!epi.begin(scratch)) return;   // same decision in every thread of every CTA

  const int tid = threadIdx.x;
  const int G = A.group;
  const int lane_g = tid & (G - 1);
  const int gid = tid / G;
  const int ngroups = OSQP_CUDA_THREADS / G;

  double asum[Epi::NSUM > 0 ? Epi::NSUM : 1];
  double amax[Epi::NMAX > 0 ? Epi::NMAX : 1];
#pragma unroll
  for (int k = 0; k < Epi::NSUM; ++k) asum[k] = 0.0;
#pragma unroll
  for (int k = 0; k < Epi::NMAX; ++k) amax[k] = -DBL_MAX;

  for (int ck = blockIdx.x; ck < A.nchunks; ck += gridDim.x) {
    const int r0 = A.chunk_row[ck], r1 = A.chunk_row[ck + 1];
    if (r0 >= r1) continue;
    const int s0 = A.rowptr[r0], s1 = A.rowptr[r1];
    int ra = r0;     // first row not yet finalised
    int par = 0;     // carry buffer parity
    for (int ts = s0;; ts += SPMV_TILE) {
      const int te = min(ts + SPMV_TILE, s1);
      const bool last = (te >= s1);
      const int a0 = ts & ~3;
      // ---- stream the tile: 4 nonzeros per thread per step, 128-bit loads
      for (int i4 = a0 + 4 * tid; i4 < te; i4 += 4 * OSQP_CUDA_THREADS) {
        const int4 c = *reinterpret_cast<const int4 *>(A.col + i4);
        const double2 v01 = *reinterpret_cast<const double2 *>(A.val + i4);
        const double2 v23 = *reinterpret_cast<const double2 *>(A.val + i4 + 2);
        int4 rr = make_int4(0, 0, 0, 0);
        if (MODE == 1) rr = *reinterpret_cast<const int4 *>(A.rowidx + i4);
        const int o = i4 - a0;
        prod[o + 0] = spmv_product<MODE>(v01.x, c.x, rr.x, g0, A.split_col);
        prod[o + 1] = spmv_product<MODE>(v01.y, c.y, rr.y, g0, A.split_col);
        prod[o + 2] = spmv_product<MODE>(v23.x, c.z, rr.z, g0, A.split_col);
        prod[o + 3] = spmv_product<MODE>(v23.y, c.w, rr.w, g0, A.split_col);
        if (NV == 2) {
          double *p1 = prod + SPMV_TILE_PAD;
          p1[o + 0] = spmv_product<0>(v01.x, c.x, 0, g1, A.split_col);
          p1[o + 1] = spmv_product<0>(v01.y, c.y, 0, g1, A.split_col);
          p1[o + 2] = spmv_product<0>(v23.x, c.z, 0, g1, A.split_col);
          p1[o + 3] = spmv_product<0>(v23.y, c.w, 0, g1, A.split_col);
        }
      }
      __syncthreads();
      // ---- rows finalised by this tile: [ra, rb)
      int rb = r1;
      if (!last) {
        // first row whose last entry lies at or beyond te (empty rows: rowptr >= te)
        int lo = ra, hi = r1;
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          const int b = A.rowptr[mid], e = A.rowptr[mid + 1];
          const bool fin = (e > b) ? (e <= te) : (b < te);
          if (fin) lo = mid + 1; else hi = mid;
        }
        rb = lo;
      }
      // the row rb is partially covered when this is not the last tile
      const int rend = last ? rb : rb + 1;
      const int nrows = rend - ra;
      const int trips = (nrows + ngroups - 1) / ngroups;
      for (int t = 0; t < trips; ++t) {
        const int r = ra + t * ngroups + gid;
        const bool valid = r < rend;
        int lo = 0, hi = 0, sp = 0;
        if (valid) {
          lo = A.rowptr[r];
          hi = A.rowptr[r + 1];
          sp = SPLIT ? A.split[r] : hi;
        }
        const int clo = max(lo, ts), chi = min(hi, te);
        double s[NACC];
#pragma unroll
        for (int k = 0; k < NACC; ++k) s[k] = 0.0;
        for (int i = clo + lane_g; i < chi; i += G) {
#pragma unroll
          for (int v = 0; v < NV; ++v) {
            const double pv = prod[v * SPMV_TILE_PAD + (i - a0)];
            if (SPLIT) {
              if (i < sp) s[v * NS] += pv; else s[v * NS + 1] += pv;
            } else {
              s[v] += pv;
            }
          }
        }
#pragma unroll
        for (int k = 0; k < NACC; ++k)
          for (int o = G >> 1; o > 0; o >>= 1) s[k] += __shfl_xor_sync(0xffffffffu, s[k], o);
        if (valid && lane_g == 0) {
          if (r == ra && lo < ts) {
#pragma unroll
            for (int k = 0; k < NACC; ++k) s[k] += carry[par * NACC + k];
          }
          if (r < rb) {
            epi.row(r, s, asum, amax);
          } else {
#pragma unroll
            for (int k = 0; k < NACC; ++k) carry[(par ^ 1) * NACC + k] = s[k];
          }
        }
      }
      __syncthreads();
      ra = rb;
      par ^= 1;
      if (last) break;
    }
  }
  // ---- per-CTA partials of the fused reductions
  if (Epi::NSUM + Epi::NMAX > 0) {
    double *out = partials + (size_t)blockIdx.x * (Epi::NSUM + Epi::NMAX);
#pragma unroll
    for (int k = 0; k < Epi::NSUM; ++k) {
      const double t = block_reduce<false>(asum[k], scratch);
      if (tid == 0) out[k] = t;
    }
#pragma unroll
    for (int k = 0; k < Epi::NMAX; ++k) {
      const double t = block_reduce<true>(amax[k], scratch);
      if (tid == 0) out[Epi::NSUM + k] = t;
    }
  }
}

template <int NV, bool SPLIT>
static inline size_t spmv_smem_bytes() {
  return (size_t)(NV * SPMV_TILE_PAD + 2 * NV * (SPLIT ? 2 : 1) + 32) * sizeof(double);
}

// Persistent grid: a multiple of the SM count, capped by the chunk count.
static inline int spmv_grid(const CsrDev &A, int ctas_per_sm) {
  int g = OSQP_CUDA_SMS * ctas_per_sm;
  if (g > OSQP_MAX_PARTIAL_BLOCKS) g = OSQP_MAX_PARTIAL_BLOCKS;
  if (g > A.nchunks) g = A.nchunks;
  if (g < 1) g = 1;
  return g;
}
