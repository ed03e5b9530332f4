// Prototype of the column-blocked SpMV as benchmarked in profiles/r2_tile_bench_v1..v4.log (host-built tiles,
// PARTIAL sums thrown away / combined by k_combine).  The product kernels are osqp_b200/csrc/spmv_tile.cuh.
// spmv_tile.cuh -- column-blocked mirror of a CSR matrix whose gathers come from SHARED memory.
//
// Replaces, for the three hot SpMVs of the ADMM iteration (w = rho .* (A p), Kp = [P | A'][p ; w],
// the rhs pass), the CSC loops of reference src/lin_alg.c:241-322 (mat_vec / mat_tpose_vec).
//
// Why: the sliced-ELL kernel (spmv_sell.cuh) issues one divergent 8-byte global gather per
// nonzero, and L1TEX serves ONE such wavefront per clock per SM whether it hits or misses
// (profiles/r2_gather_rate2_microbench.log: 0.96 gathers/clk/SM from an L2-resident table);
// that pins both CG SpMVs at ~0.4 of the HBM roofline.  Shared memory serves random fp64
// reads at 4.75 per clock per SM (same log), and the SpMV inner loop with the table in shared
// memory streams its (col, val) pairs at 5.2 TB/s.  Distributed shared memory does not
// (0.35 gathers/clk/SM through ld.shared::cluster), so the table must be LOCAL to the CTA:
//   * the columns are cut into blocks of `cb` <= TILE_CB_MAX entries (<= 216 KB of shared memory);
//   * block c of the matrix -- all rows, only the entries whose column lies in the block -- is cut
//     into WINDOWS of 1024 consecutive rows.  Inside a window the rows are sorted by their entry
//     count in this block (stable), so that a SLICE of 32 sorted rows (lane l IS sorted row
//     32 s + l) has nearly uniform length: slices are stored pass-major, padded to their longest
//     row (col = 0, val = 0; ~5 % padding on binomial counts instead of the 2.4x of unsorted
//     slices), with 16-bit local column indices -- 10 bytes per stored entry instead of 12;
//   * a CTA works through ITEMS: (block, run of slices).  Per item it pulls the block's part of
//     the gathered vector into shared memory with TMA bulk copies (cp.async.bulk + mbarrier
//     complete_tx); then every warp free-runs over its own contiguous range of the run (ranges
//     balanced by passes at setup) the way the sliced-ELL kernel does -- one contiguous stream
//     per warp, chunks of TILE_CH passes, the next chunk's col/val in flight while the current
//     one is consumed, no CTA-wide barrier on the row path -- and writes the partial row sums in
//     SORTED order, coalesced, to partial[c][window * 1024 + sorted position];
//   * the kernel that runs the fused row epilogue (k_tile_epi in engine.cu) takes a window per CTA,
//     stages the window's partial sums of a group of blocks in shared memory, lets every row pick
//     its own through inv[c][row] and adds them in block order: fixed order -> bit-reproducible.
// Pays when a row keeps >= ~2 entries per block it touches (config 2: A has 10 entries per row
// in 4 blocks, A' 100 per row in 37-38); otherwise the partial sums outweigh the matrix and the
// sliced-ELL kernel stays in charge (tile_worth in dev_mats.cu).
#pragma once
#include "../../osqp_b200/csrc/spmv_common.cuh"

#define TILE_THREADS 1024      // one CTA per SM (the table takes the shared memory)
#define TILE_WARPS 32
#define TILE_WIN 1024          // rows per sorting window
#ifndef TILE_CH
#define TILE_CH 4              // passes (of 32 entries) per chunk
#endif
#define TILE_CB_MAX 27648      // doubles of shared memory for the table (216 KB)
#ifndef TILE_LD
#define TILE_LD 0              // col/val loads: 0 evict-first (ld.global.cs), 1 L1::no_allocate
#endif
__device__ __forceinline__ unsigned tile_ld_col(const unsigned short *p) {
#if TILE_LD == 1
  unsigned short v; asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(v) : "l"(p)); return v;
#else
  return __ldcs(p);
#endif
}
__device__ __forceinline__ double tile_ld_val(const double *p) {
#if TILE_LD == 1
  double v; asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p)); return v;
#else
  return __ldcs(p);
#endif
}

struct TileView {
  int rows, rows_pad, cols, cb, nblocks, nwin, ncta;
  const int *slice_ptr;         // [nblocks * nwin * 32 + 1]: first pass (of 32 entries) of the slice
  const unsigned short *inv;    // [nblocks][rows_pad]: row -> its sorted position inside the window
  const unsigned short *col;    // [passes * 32]: column minus c * cb
  const double *val;            // [passes * 32]
  const int *item_ptr;          // [ncta + 1]: items of every CTA
  const int *item_blk;          // [nitems]: column block of the item
  const int *item_ws;           // [nitems][33]: first slice of every warp (global slice index)
};
static inline size_t tile_smem_bytes(const TileView &T, int nv) { return (size_t)T.cb * 8 * nv; }

template <int NV>
__global__ void __launch_bounds__(TILE_THREADS, 1)
spmv_tile_kernel(TileView T, const double *__restrict__ x0, const double *__restrict__ x1,
                 double *__restrict__ partial, const int *skip) {
  constexpr int CH = TILE_CH;
  extern __shared__ __align__(128) double tile_smem[];
  __shared__ unsigned long long bar;
  if (skip && *skip) return;       // CG already converged: the launch is a no-op
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double *sx0 = tile_smem, *sx1 = tile_smem + ((NV == 2) ? T.cb : 0);
  const unsigned sb0 = smem_u32(sx0), sb1 = smem_u32(sx1);
  if (tid == 0) {
    mbar_init(&bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  const int it_begin = __ldg(T.item_ptr + blockIdx.x), it_end = __ldg(T.item_ptr + blockIdx.x + 1);
  int tphase = 0;
  for (int it = it_begin; it < it_end; ++it) {
    const int blk = __ldg(T.item_blk + it);
    __syncthreads();                                // mbarrier ready / nobody gathers from the old table
    {
      // ---- the block's part of the gathered vector(s) -> shared memory (TMA bulk copies)
      const int c0 = blk * T.cb;
      const int ncols = min(T.cb, T.cols - c0);
      if (tid == 0) {
        fence_proxy_async();
        const unsigned even = (unsigned)(ncols & ~1);   // bulk copies move multiples of 16 bytes
        mbar_expect_tx(&bar, even * 8u * NV);
        for (unsigned o = 0; o < even; o += 4096u) {
          const unsigned len = min(4096u, even - o);
          bulk_g2s(sx0 + o, x0 + c0 + o, len * 8u, &bar);
          if (NV == 2) bulk_g2s(sx1 + o, x1 + c0 + o, len * 8u, &bar);
        }
        if (ncols & 1) {
          sx0[ncols - 1] = x0[c0 + ncols - 1];
          if (NV == 2) sx1[ncols - 1] = x1[c0 + ncols - 1];
        }
      }
    }
    // ---- this warp's contiguous range of slices [s, s_end) = passes [P0, P1)
    int s = __ldg(T.item_ws + it * 33 + warp);
    const int s_end = __ldg(T.item_ws + it * 33 + warp + 1);
    const int P0 = __ldg(T.slice_ptr + s), P1 = __ldg(T.slice_ptr + s_end);
    const int nch = (P1 - P0 + CH - 1) / CH;
    unsigned cn[CH]; double vn[CH];
    auto load = [&](int ch) {
      const int np = min(CH, P1 - P0 - ch * CH);
      const size_t e0 = ((size_t)(P0 + ch * CH)) * 32 + lane;
#pragma unroll
      for (int j = 0; j < CH; ++j) {
        cn[j] = 0; vn[j] = 0.0;
        if (j < np) { cn[j] = tile_ld_col(T.col + e0 + j * 32); vn[j] = tile_ld_val(T.val + e0 + j * 32); }
      }
    };
    if (nch > 0) load(0);                           // streams in while the table lands
    int pc = P0;
    int p1 = (s < s_end) ? __ldg(T.slice_ptr + s + 1) : P0;
    int p2 = (s + 1 < s_end) ? __ldg(T.slice_ptr + s + 2) : p1;
    int left = p1 - pc;
    double *out = partial;      // slices are numbered globally: slot (s * 32 + lane) = blk * rows_pad + sorted row
    double a0 = 0.0, a1 = 0.0;
    auto finish = [&]() {                           // close slice s, open slice s + 1
      const size_t r = (size_t)s * 32 + lane;
      if (NV == 1) out[r] = a0; else { out[2 * r] = a0; out[2 * r + 1] = a1; }
      a0 = 0.0; a1 = 0.0;
      ++s;
      pc = p1; p1 = p2;
      if (s + 1 < s_end) p2 = __ldg(T.slice_ptr + s + 2);
      left = p1 - pc;
    };
    // empty slices (rows without entries in this block) are never written: partial[] is zeroed
    // once at setup and nothing else touches their slots
    while (s < s_end && left == 0) { ++s; pc = p1; p1 = p2; if (s + 1 < s_end) p2 = __ldg(T.slice_ptr + s + 2); left = p1 - pc; }
    __syncthreads();                                // the odd tail element is visible
    mbar_wait(&bar, tphase);                        // the table has landed
    tphase ^= 1;
    for (int ch = 0; ch < nch; ++ch) {
      const int np = min(CH, P1 - P0 - ch * CH);
      unsigned c[CH]; double v[CH], g0[CH], g1[NV == 2 ? CH : 1];
#pragma unroll
      for (int j = 0; j < CH; ++j) { c[j] = cn[j]; v[j] = vn[j]; }
      if (ch + 1 < nch) load(ch + 1);
#pragma unroll
      for (int j = 0; j < CH; ++j) {
        asm volatile("ld.shared.f64 %0, [%1];" : "=d"(g0[j]) : "r"(sb0 + 8u * c[j]));
        if (NV == 2) asm volatile("ld.shared.f64 %0, [%1];" : "=d"(g1[j]) : "r"(sb1 + 8u * c[j]));
      }
#pragma unroll
      for (int j = 0; j < CH; ++j) {
        if (j < np) {
          a0 = fma(v[j], g0[j], a0);
          if (NV == 2) a1 = fma(v[j], g1[j], a1);
          if (--left == 0) {
            finish();
            while (s < s_end && left == 0) { ++s; pc = p1; p1 = p2; if (s + 1 < s_end) p2 = __ldg(T.slice_ptr + s + 2); left = p1 - pc; }
          }
        }
      }
    }
  }
}
