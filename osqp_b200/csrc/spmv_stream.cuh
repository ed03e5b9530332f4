// spmv_stream.cuh -- row-wise CSR SpMV for B200, "stream through shared memory".
//
// Replaces the CSC scatter/gather loops of reference src/lin_alg.c:241-322
// (mat_vec / mat_tpose_vec) and, through the row epilogues, the vector loops of
// src/auxil.c:161-225 that consume their results.
//
// Design (HBM / L1TEX-gather bound, no tensor cores):
//  * the nonzeros of a CSR matrix are cut into CHUNKS of whole rows: chunk k owns
//    the rows whose first nonzero lies in [k*CHUNK_NNZ, (k+1)*CHUNK_NNZ).  A
//    persistent grid (a multiple of the 148 SMs) walks the chunks round-robin,
//    so the work split depends on the matrix only -> bit-reproducible sums;
//  * val[] / col[] are contiguous streams: each CTA copies the NEXT tile into a
//    second shared-memory buffer with 16-byte cp.async (LDGSTS, L1-bypassing)
//    while it works on the current one, so HBM latency is off the critical path;
//  * every thread then issues its 8 gathers of the dense vector back to back
//    (independent LDGs, L2 hits), multiplies and parks the products in place of
//    the values;
//  * groups of G lanes (G = 1..32 chosen from the mean row length: "vector per
//    row" / "warp per row") reduce each row segment out of shared memory in a
//    fixed order and run a fused row epilogue (rho scaling, ADMM updates,
//    residual norms, dot-product partials ...);
//  * rows longer than a tile are carried across tiles inside the CTA.
#pragma once
#include "dev_common.cuh"

#define SPMV_TILE      2048   // nonzeros staged per tile
#define SPMV_CHUNK_NNZ 2048   // chunk quantum; a chunk spans < CHUNK + longest row
#define SPMV_TILE_PAD  (SPMV_TILE + 8)
#define SPMV_PER_THREAD (SPMV_TILE / OSQP_CUDA_THREADS)

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

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// One unit of work of a CTA: tile [ts, te) of chunk ck (rows [r0, r1), entries [s0, s1)).
struct SpmvItem {
  int ck, r0, r1, s0, s1, ts;
  bool valid;
};
__device__ __forceinline__ void spmv_item_load_chunk(SpmvItem &it, const CsrView &A) {
  it.valid = false;
  while (it.ck < A.nchunks) {
    it.r0 = A.chunk_row[it.ck];
    it.r1 = A.chunk_row[it.ck + 1];
    if (it.r0 < it.r1) {
      it.s0 = A.rowptr[it.r0];
      it.s1 = A.rowptr[it.r1];
      it.ts = it.s0;
      it.valid = true;
      return;
    }
    it.ck += gridDim.x;
  }
}
__device__ __forceinline__ void spmv_item_next(SpmvItem &it, const CsrView &A) {
  if (it.ts + SPMV_TILE < it.s1) { it.ts += SPMV_TILE; return; }
  it.ck += gridDim.x;
  spmv_item_load_chunk(it, A);
}
// cp.async the tile of `it` into (scol, sval): 16-byte pieces, coalesced
__device__ __forceinline__ void spmv_prefetch(const SpmvItem &it, const CsrView &A, int *scol,
                                              double *sval) {
  if (it.valid) {
    const int te = min(it.ts + SPMV_TILE, it.s1);
    const int a0 = it.ts & ~3;
    const int cnt = te - a0;                       // <= TILE + 3
    for (int k = 4 * threadIdx.x; k < cnt; k += 4 * OSQP_CUDA_THREADS) cp_async16(scol + k, A.col + a0 + k);
    for (int k = 2 * threadIdx.x; k < cnt; k += 2 * OSQP_CUDA_THREADS) cp_async16(sval + k, A.val + a0 + k);
  }
  cp_async_commit();
}

// Epilogue concept:
//   static constexpr int NSUM, NMAX;          per-thread accumulators reduced per CTA
//   __device__ bool begin(double *scratch);   CTA-uniform prologue; false = skip kernel
//   __device__ void row(int r, const double *s, double *asum, double *amax) const;
// where s[v * NS + part] is the row sum of gathered vector v (part 1 = A' part
// when SPLIT).  Partials land in partials[blockIdx.x * (NSUM + NMAX) + k].
// non-blocking gather: the load is issued here, the value is consumed later
__device__ __forceinline__ double ldg_issue(const double *p) {
  double v;
  asm volatile("ld.global.nc.f64 %0, [%1];\n" : "=d"(v) : "l"(p) : "memory");
  return v;
}

template <int NV, bool SPLIT, int MODE, class Epi>
__global__ void __launch_bounds__(OSQP_CUDA_THREADS, (NV == 2 ? 2 : 3))
spmv_stream_kernel(CsrView A, Gather g0, Gather g1, Epi epi_in, double *partials) {
  constexpr int NS = SPLIT ? 2 : 1;
  constexpr int NACC = NV * NS;
  constexpr int SLOTS = SPMV_PER_THREAD + 1;   // a tile holds up to TILE + 3 entries
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double *sval0 = reinterpret_cast<double *>(smem_raw);                   // 3 x TILE_PAD doubles
  double *prodB = sval0 + 3 * SPMV_TILE_PAD;                              // NV == 2: 2 x TILE_PAD doubles
  double *carry = prodB + (NV == 2 ? 2 * SPMV_TILE_PAD : 0);              // 2 * NACC
  double *scratch = carry + 2 * NACC;                                     // 32
  int *scol0 = reinterpret_cast<int *>(scratch + 32);                     // 3 x TILE_PAD ints

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

  // Three shared-memory tile buffers rotate through three roles:
  //   b0: products of tile `cur` (being row-reduced)
  //   b1: col/val of tile `nxt` (landed; its gathers are in flight during the reduction)
  //   b2: cp.async target for the tile after `nxt`
  SpmvItem cur, nxt, nn;
  cur.ck = blockIdx.x;
  spmv_item_load_chunk(cur, A);
  int b0 = 0, b1 = 1, b2 = 2;
  spmv_prefetch(cur, A, scol0 + b0 * SPMV_TILE_PAD, sval0 + b0 * SPMV_TILE_PAD);
  nxt = cur;
  if (cur.valid) spmv_item_next(nxt, A);
  spmv_prefetch(nxt, A, scol0 + b1 * SPMV_TILE_PAD, sval0 + b1 * SPMV_TILE_PAD);

  double x[SLOTS], y[NV == 2 ? SLOTS : 1];

  // gathers of one landed tile: issue (registers x / y hold the pending loads) ...
  auto issue = [&](const SpmvItem &it, int b) {
    const int te = min(it.ts + SPMV_TILE, it.s1);
    const int a0 = it.ts & ~3;
    const int cnt = te - a0;
    const int *scol = scol0 + b * SPMV_TILE_PAD;
#pragma unroll
    for (int k = 0; k < SLOTS; ++k) {
      const int o = tid + k * OSQP_CUDA_THREADS;
      x[k] = 0.0;
      if (NV == 2) y[k] = 0.0;
      if (o < cnt) {
        const int c = scol[o];
        if (MODE == 0) {
          x[k] = ldg_issue((c < A.split_col) ? g0.lo + c : g0.hi + (c - A.split_col));
        } else {
          // Jacobi: the P part keeps its diagonal entry (marker -1), the A' part val^2 * rho
          if (c < A.split_col) x[k] = (c == A.rowidx[a0 + o]) ? -1.0 : 0.0;
          else x[k] = ldg_issue(g0.hi + (c - A.split_col));
        }
        if (NV == 2) y[k] = ldg_issue((c < A.split_col) ? g1.lo + c : g1.hi + (c - A.split_col));
      }
    }
  };
  // ... and complete: the products overwrite the values in place
  auto complete = [&](const SpmvItem &it, int b) {
    const int te = min(it.ts + SPMV_TILE, it.s1);
    const int a0 = it.ts & ~3;
    const int cnt = te - a0;
    double *prod = sval0 + b * SPMV_TILE_PAD;
    double *prod1 = prodB + (b & 1) * SPMV_TILE_PAD;
    const int *scol = scol0 + b * SPMV_TILE_PAD;
#pragma unroll
    for (int k = 0; k < SLOTS; ++k) {
      const int o = tid + k * OSQP_CUDA_THREADS;
      if (o < cnt) {
        const double v = prod[o];
        if (MODE == 0) {
          prod[o] = v * x[k];
        } else {
          prod[o] = (scol[o] < A.split_col) ? (x[k] < 0.0 ? v : 0.0) : v * v * x[k];
        }
        if (NV == 2) prod1[o] = v * y[k];
      }
    }
  };

  if (cur.valid) {
    cp_async_wait<1>();
    __syncthreads();
    issue(cur, b0);
    complete(cur, b0);
    __syncthreads();
  }
  int ra = cur.r0;   // first row of the chunk not yet finalised
  int par = 0;       // carry buffer parity
  int pb0 = 0;       // which half of prodB holds the second-vector products of `cur`

  while (cur.valid) {
    nn = nxt;
    if (nxt.valid) spmv_item_next(nn, A);
    spmv_prefetch(nn, A, scol0 + b2 * SPMV_TILE_PAD, sval0 + b2 * SPMV_TILE_PAD);
    cp_async_wait<1>();          // tile `nxt` has landed (per thread) ...
    __syncthreads();             // ... and is visible to the CTA
    if (nxt.valid) issue(nxt, b1);

    // ---- row reduction of `cur` out of b0 while the gathers of `nxt` are in flight
    const double *prod = sval0 + b0 * SPMV_TILE_PAD;
    const double *prod1 = prodB + pb0 * SPMV_TILE_PAD;
    const int ts = cur.ts, s1 = cur.s1, r1 = cur.r1;
    const int te = min(ts + SPMV_TILE, s1);
    const bool last = (te >= s1);
    const int a0 = ts & ~3;
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
        const double pv = prod[i - a0];
        if (SPLIT) {
          if (i < sp) s[0] += pv; else s[1] += pv;
        } else {
          s[0] += pv;
        }
        if (NV == 2) {
          const double pw = prod1[i - a0];
          if (SPLIT) {
            if (i < sp) s[NS] += pw; else s[NS + 1] += pw;
          } else {
            s[NS] += pw;
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
    // ---- products of `nxt` (its gathers have had the whole reduction to arrive)
    if (nxt.valid) complete(nxt, b1);
    __syncthreads();
    // ---- rotate
    if (nxt.valid && nxt.ck == cur.ck) { ra = rb; par ^= 1; }
    else { ra = nxt.r0; par = 0; }
    cur = nxt;
    nxt = nn;
    pb0 = b1 & 1;
    const int t0 = b0; b0 = b1; b1 = b2; b2 = t0;
  }
  cp_async_wait<0>();
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
  return (size_t)((3 + (NV == 2 ? 2 : 0)) * SPMV_TILE_PAD + 2 * NV * (SPLIT ? 2 : 1) + 32) * sizeof(double) +
         (size_t)3 * SPMV_TILE_PAD * sizeof(int);
}

// Persistent grid: a multiple of the SM count, capped by the chunk count.
static inline int spmv_grid(const CsrDev &A, int ctas_per_sm) {
  int g = OSQP_CUDA_SMS * ctas_per_sm;
  if (g > OSQP_MAX_PARTIAL_BLOCKS) g = OSQP_MAX_PARTIAL_BLOCKS;
  if (g > A.nchunks) g = A.nchunks;
  if (g < 1) g = 1;
  return g;
}
