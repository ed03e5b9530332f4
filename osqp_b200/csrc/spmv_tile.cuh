// spmv_tile.cuh -- column-blocked mirrors of a CSR matrix whose gathers come from SHARED memory.
//
// Replaces, for the hot SpMVs of the ADMM iteration (w = rho .* (A p), Kp = [P | A'][p ; w], the
// rhs pass), the CSC loops of reference src/lin_alg.c:241-322 (mat_vec / mat_tpose_vec).
//
// Why: the sliced-ELL kernel (spmv_sell.cuh) issues one divergent 8-byte global gather per
// nonzero, and L1TEX serves ONE such wavefront per clock per SM whether it hits or misses
// (profiles/r2_gather_rate2_microbench.log: 0.96 gathers/clk/SM from an L2-resident table);
// that pins both CG SpMVs at ~0.4 of the HBM roofline.  Shared memory serves random fp64
// reads at 4.75 per clock per SM (same log), and the SpMV inner loop with the table in shared
// memory streams its (col, val) pairs at 5.2 TB/s.  Distributed shared memory does not
// (0.35 gathers/clk/SM through ld.shared::cluster), so the table must be LOCAL to the CTA.
//
// Common layout.  The columns are cut into blocks of `cb` entries.  Block c of the matrix -- all
// rows, only the entries whose column lies in the block -- is cut into WINDOWS of `win`
// consecutive rows.  Inside a window the rows are sorted by their entry count in this block
// (stable), so that a SLICE of 32 sorted rows (lane l IS sorted row 32 k + l of the window) has
// nearly uniform length: slices are stored pass-major, padded to their longest row (col = 0,
// val = 0; 5-8 % padding on binomial counts instead of the 2.4x of unsorted slices), with 16-bit
// local column indices -- 10 bytes per stored entry instead of 12.  Slices are numbered block-major:
// slice (c * nwin + window) * (win / 32) + k; slice_ptr[] holds their first pass.  A warp always
// walks a CONTIGUOUS range of slices = one contiguous run of the col and val streams, in chunks of
// TILE_CH passes, the next chunk in flight while the current one is consumed (as spmv_sell.cuh).
// The block's part of the gathered vector arrives in shared memory by TMA bulk copies
// (cp.async.bulk + mbarrier complete_tx).
//
// Two kernels, by where the row sums can live:
//  * ROWS mode (spmv_tile_rows_kernel; A: many short rows, few column blocks).  A CTA owns a range
//    of windows for ALL column blocks: it walks block 0, 1, ... of its rows, reloading the table
//    in between, and accumulates the row sums in SHARED memory (un-sorting on the way: sorted
//    position -> row through perm[]).  After the last block it runs the fused row epilogue on its
//    rows, coalesced.  No partial sums ever reach memory; each row sum is added up in block order
//    by one lane at a time -> bit-reproducible.  Costs every CTA a read of the whole gathered
//    vector from L2 (148 x 800 KB on BASELINE config 2, against a 100 MB matrix stream).
//  * PARTIAL mode (spmv_tile_kernel + k_tile_epi; the A' part of M = [P | A']: the gathered
//    m-vector is 8 MB, so a CTA can only afford ONE block of it).  A CTA works through items
//    (block, run of slices) and writes partial row sums in SORTED order, coalesced, to
//    partial[c][window * win + sorted position]; k_tile_epi takes a window per CTA, stages the
//    partial sums of a group of blocks in shared memory, lets every row pick its own through
//    inv[c][row], adds them in block order (bit-reproducible) and runs the fused row epilogue.
// Pays when a row keeps >= ~1.5 entries per block it touches; otherwise the sliced-ELL kernel
// stays in charge (tile_build in dev_mats.cu).
#pragma once
#include "spmv_common.cuh"

#define TILE_THREADS 1024      // one CTA per SM (the table takes the shared memory)
#define TILE_WARPS 32
#ifndef TILE_CH
#define TILE_CH 4              // passes (of 32 entries) per chunk
#endif
#define TILE_ROWS_ITEMS 9      // ROWS mode: a CTA owns (and the build sorts) at most 1024 * 9 rows
#define TILE_ROWS_MAX (1024 * TILE_ROWS_ITEMS)
#ifndef TILE_WIN_PART
#define TILE_WIN_PART 512      // rows per sorting window (256: product 35.5 -> 38.3 us, combine 18.3 -> 15.5: no gain), PARTIAL mode (a window = one k_tile_epi CTA)
#endif
#define TEPI_CTAS_PER_SM (1024 / TILE_WIN_PART)
#define TILE_SMEM_MAX (224 * 1024)   // dynamic shared memory the kernels may ask for

struct TileView {
  int rows, rows_pad, cols, cb, nblocks, nwin, win, ncta, racc;
  int n_lo, cb_lo, nb_lo;       // first column range (gathered from xlo): nb_lo blocks of cb_lo; the others from xhi
  int nsb;                      // slices per block (ROWS: sum over the CTAs of ceil(rows of the CTA / 32))
  const int *slice_ptr;         // [nblocks * nsb + 1]: first pass (of 32 entries) of the slice
  const unsigned short *inv;    // PARTIAL: [nblocks][rows_pad] row -> sorted position inside its window
  const unsigned short *perm;   // ROWS:    [nblocks][nsb * 32] sorted position -> row inside its CTA's range
  const unsigned short *col;    // [passes * 32]: column minus c * cb
  const double *val;            // [passes * 32]
  const int *item_ptr;          // PARTIAL: [ncta + 1] items of every CTA
  const int *item_blk;          // PARTIAL: [nitems] column block of the item
  const int *item_ws;           // PARTIAL: [nitems][33] first non-empty-slice index (into ne_*) of every warp;  ROWS: [ncta][nblocks][33]
  const int *cta_row;           // ROWS: [ncta + 1] first row of every CTA
  const int *cta_soff;          // ROWS: [ncta + 1] first slice (inside a block) of every CTA
  const int2 *ne;               // [K + 1]: the K slices that have entries, ascending: {first pass, slice number}; then {passes, -1}
  const int *wb_ptr, *wb_blk;   // PARTIAL: [nwin + 1], [..] ascending column blocks in which the window has entries
};
// block -> its part of the gathered vectors
__device__ __forceinline__ const double *tile_block_src(const TileView &T, int blk, const double *xlo, const double *xhi,
                                                        int &ncols) {
  if (blk < T.nb_lo) { const int c0 = blk * T.cb_lo; ncols = min(T.cb_lo, T.n_lo - c0); return xlo + c0; }
  const int c0 = (blk - T.nb_lo) * T.cb;
  ncols = min(T.cb, T.cols - c0);
  return xhi + c0;
}

// ---- shared by both kernels: TMA load of one block of the gathered vector into `sx`
// (+ `xbytes` of `xsrc` into `xdst`, a multiple of 16: the ROWS kernel's perm slab)
__device__ __forceinline__ void tile_load_table(double *sx, const double *x, int ncols,
                                                unsigned long long *bar, void *xdst = nullptr,
                                                const void *xsrc = nullptr, unsigned xbytes = 0) {
  fence_proxy_async();                              // generic-proxy reads of the old table precede the async writes
  const unsigned even = (unsigned)(ncols & ~1);     // bulk copies move multiples of 16 bytes
  mbar_expect_tx(bar, even * 8u + xbytes);
  for (unsigned o = 0; o < xbytes; o += 32768u)
    bulk_g2s((char *)xdst + o, (const char *)xsrc + o, min(32768u, xbytes - o), bar);
  for (unsigned o = 0; o < even; o += 4096u) {
    const unsigned len = min(4096u, even - o);
    bulk_g2s(sx + o, x + o, len * 8u, bar);
  }
  if (ncols & 1) sx[ncols - 1] = x[ncols - 1];
}

// ---- shared by both kernels: one warp walks entries [k, k_end) of the list of NON-EMPTY slices
// (T.ne: structured matrices leave most slices of a block empty, and those must cost
// nothing), gathering from the table at shared address `sb`; done(slice, sum) is called once per
// slice, in order.  `wait` is called once, after the first chunk's loads are issued and before the
// first gather (the table lands meanwhile).
template <class Wait, class Done>
__device__ __forceinline__ void tile_walk(const TileView &T, int k, const int k_end, const unsigned sb,
                                          Wait wait, Done done) {
  constexpr int CH = TILE_CH;
  const int lane = threadIdx.x & 31;
  const int2 e0 = __ldg(T.ne + k);
  const int P0 = e0.x, P1 = __ldg(&T.ne[k_end].x);
  const int nch = (P1 - P0 + CH - 1) / CH;
  unsigned cn[CH]; double vn[CH];
  auto load = [&](int ch) {
    const int np = min(CH, P1 - P0 - ch * CH);
    const size_t e0 = ((size_t)(P0 + ch * CH)) * 32 + lane;
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      cn[j] = 0; vn[j] = 0.0;
      if (j < np) { cn[j] = __ldcs(T.col + e0 + j * 32); vn[j] = __ldcs(T.val + e0 + j * 32); }
    }
  };
  if (nch > 0) load(0);                             // streams in while the table lands
  int pc = P0, sid = e0.y;
  int2 e1 = (k < k_end) ? __ldg(T.ne + k + 1) : e0;       // {end of the current slice, number of the next one}
  int2 e2 = (k + 1 < k_end) ? __ldg(T.ne + k + 2) : e1;   // one more ahead: a slice can be a single pass
  int left = e1.x - pc;
  auto next = [&]() {
    ++k; pc = e1.x; sid = e1.y; e1 = e2;
    if (k + 1 < k_end) e2 = __ldg(T.ne + k + 2);
    left = e1.x - pc;
  };
  wait();
  double a0 = 0.0;
  for (int ch = 0; ch < nch; ++ch) {
    const int np = min(CH, P1 - P0 - ch * CH);
    unsigned c[CH]; double v[CH], g[CH];
#pragma unroll
    for (int j = 0; j < CH; ++j) { c[j] = cn[j]; v[j] = vn[j]; }
    if (ch + 1 < nch) load(ch + 1);
#pragma unroll
    for (int j = 0; j < CH; ++j) asm volatile("ld.shared.f64 %0, [%1];" : "=d"(g[j]) : "r"(sb + 8u * c[j]));
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      if (j < np) {
        a0 = fma(v[j], g[j], a0);
        if (--left == 0) {
          done(sid, a0);
          a0 = 0.0;
          next();
        }
      }
    }
  }
}

// =============================================================================== ROWS mode
// The sorting "window" of a CTA is its whole row range (<= TILE_ROWS_MAX rows, balanced by nonzeros
// at setup): slices of a block are numbered CTA by CTA, nsb per block.
// Dynamic shared memory: table[cb], acc[racc] (racc = slices of the largest CTA * 32), then the
// CTA's slab of perm[c] for the current block (racc x u16; arrives with the table).
template <class Epi>
__global__ void __launch_bounds__(TILE_THREADS, 1)
spmv_tile_rows_kernel(TileView T, const double *__restrict__ xlo, const double *__restrict__ xhi, Epi epi_in,
                      double *partials) {
  extern __shared__ __align__(128) double tile_smem[];
  __shared__ unsigned long long bar;
  __shared__ double scratch[32];
  __shared__ unsigned sflag[128];
  Epi epi = epi_in;
  if (!epi.begin(scratch)) return;                  // same decision in every thread of every CTA
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double *sx = tile_smem, *acc = tile_smem + max(T.cb, T.nb_lo ? T.cb_lo : 0);
  unsigned short *sperm = reinterpret_cast<unsigned short *>(acc + T.racc);
  const unsigned sb = smem_u32(sx), sa = smem_u32(acc), sp = smem_u32(sperm);
  const int r0 = __ldg(T.cta_row + blockIdx.x), nrows = __ldg(T.cta_row + blockIdx.x + 1) - r0;
  const int so = __ldg(T.cta_soff + blockIdx.x), ns = __ldg(T.cta_soff + blockIdx.x + 1) - so;
  if (tid == 0) {
    mbar_init(&bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  for (int i = tid; i < ns * 32; i += TILE_THREADS) acc[i] = 0.0;
  // blocks in which none of this CTA's rows has an entry cost nothing: matrices whose column range is
  // much wider than what a row range touches (Lasso: [A_d  -I  0]).  One bit per block, all tests in
  // flight at once (nblocks <= 4096).
  for (int c0 = 0; c0 < T.nblocks; c0 += TILE_THREADS) {
    const int c = c0 + tid;
    const bool has = c < T.nblocks &&
                     __ldg(T.slice_ptr + (size_t)c * T.nsb + so) != __ldg(T.slice_ptr + (size_t)c * T.nsb + so + ns);
    const unsigned bits = __ballot_sync(0xffffffffu, has);
    if (lane == 0) sflag[(c0 >> 5) + warp] = bits;
  }
  __syncthreads();
  unsigned tphase = 0;
  for (int c = 0; c < T.nblocks; ++c) {
    if (!((sflag[c >> 5] >> (c & 31)) & 1u)) continue;
    __syncthreads();                                // acc zeroed / mbarrier ready / old table no longer read
    if (tid == 0) {
      int ncols;
      const double *src = tile_block_src(T, c, xlo, xhi, ncols);
      tile_load_table(sx, src, ncols, &bar, sperm, T.perm + ((size_t)c * T.nsb + so) * 32, (unsigned)(ns * 64));
    }
    const int *ws = T.item_ws + ((size_t)blockIdx.x * T.nblocks + c) * 33;
    const int s_lo = __ldg(ws + warp), s_hi = __ldg(ws + warp + 1);
    const int sbase = c * T.nsb + so;               // first slice of block c that this CTA owns
    tile_walk(T, s_lo, s_hi, sb,
              [&]() { __syncthreads(); mbar_wait(&bar, tphase); },
              [&](int s, double sum) {
                // slice s = sorted rows [32 k, 32 k + 32) of this CTA's range; un-sort into acc
                const unsigned o = (unsigned)(s - sbase) * 32u + lane;
                unsigned short pr;
                asm volatile("ld.shared.u16 %0, [%1];" : "=h"(pr) : "r"(sp + 2u * o));
                double t;
                asm volatile("ld.shared.f64 %0, [%1];" : "=d"(t) : "r"(sa + 8u * pr));
                t += sum;
                asm volatile("st.shared.f64 [%0], %1;" ::"r"(sa + 8u * pr), "d"(t) : "memory");
              });
    tphase ^= 1;
  }
  __syncthreads();
  // ---- fused row epilogue over this CTA's rows, coalesced
  double asum[Epi::NSUM > 0 ? Epi::NSUM : 1];
  double amax[Epi::NMAX > 0 ? Epi::NMAX : 1];
#pragma unroll
  for (int k = 0; k < Epi::NSUM; ++k) asum[k] = 0.0;
#pragma unroll
  for (int k = 0; k < Epi::NMAX; ++k) amax[k] = -DBL_MAX;
  for (int i = tid; i < nrows; i += TILE_THREADS) {
    const typename Epi::Pre pre = epi.load(r0 + i);
    const double sum = acc[i];
    epi.row(r0 + i, &sum, pre, asum, amax);
  }
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

// ============================================================================ PARTIAL mode
// Epi: only its CTA-uniform prologue begin() is used here (it can veto the launch: CG iterations
// enqueued beyond convergence); the rows' epilogue runs in k_tile_epi, which calls begin() again,
// so begin() must be idempotent (epi_tile_ok lists the epilogues for which it is).
// Dynamic shared memory: table[cb].
template <class Epi>
__global__ void __launch_bounds__(TILE_THREADS, 1)
spmv_tile_kernel(TileView T, const double *__restrict__ xlo, const double *__restrict__ xhi,
                 double *__restrict__ partial, Epi epi_in) {
  extern __shared__ __align__(128) double tile_smem[];
  __shared__ unsigned long long bar;
  __shared__ double scratch[32];
  {
    Epi epi = epi_in;
    if (!epi.begin(scratch)) return;                // same decision in every thread of every CTA
  }
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned sb = smem_u32(tile_smem);
  if (tid == 0) {
    mbar_init(&bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  const int it_begin = __ldg(T.item_ptr + blockIdx.x), it_end = __ldg(T.item_ptr + blockIdx.x + 1);
  unsigned tphase = 0;
  for (int it = it_begin; it < it_end; ++it) {
    const int blk = __ldg(T.item_blk + it);
    __syncthreads();                                // mbarrier ready / nobody gathers from the old table
    if (tid == 0) {
      int ncols;
      const double *src = tile_block_src(T, blk, xlo, xhi, ncols);
      tile_load_table(tile_smem, src, ncols, &bar);
    }
    const int s_lo = __ldg(T.item_ws + it * 33 + warp), s_hi = __ldg(T.item_ws + it * 33 + warp + 1);
    // slices are numbered block-major, so slot (s * 32 + lane) IS partial[blk][window * win + sorted
    // position]; slices without entries are never written: partial[] is zeroed once at setup
    tile_walk(T, s_lo, s_hi, sb,
              [&]() { __syncthreads(); mbar_wait(&bar, tphase); },
              [&](int s, double sum) { partial[(size_t)s * 32 + lane] = sum; });
    tphase ^= 1;
  }
}

// Row epilogue over the partial sums: one CTA of `win` = 512 threads per window (persistent grid),
// one row per thread.  The window's partial sums of TEPI_G blocks at a time are staged in shared
// memory, coalesced; every row picks its own through inv[c][row] and adds them in block order; the
// next group's loads are in flight meanwhile.
#define TEPI_G 8
template <class Epi>
__global__ void __launch_bounds__(TILE_WIN_PART, TEPI_CTAS_PER_SM)
k_tile_epi(Epi epi_in, TileView T, const double *__restrict__ partial, double *partials_out) {
  __shared__ double sp[TEPI_G][TILE_WIN_PART];
  __shared__ double scratch[32];
  Epi epi = epi_in;
  if (!epi.begin(scratch)) return;
  const int tid = threadIdx.x;
  double asum[Epi::NSUM > 0 ? Epi::NSUM : 1];
  double amax[Epi::NMAX > 0 ? Epi::NMAX : 1];
#pragma unroll
  for (int k = 0; k < Epi::NSUM; ++k) asum[k] = 0.0;
#pragma unroll
  for (int k = 0; k < Epi::NMAX; ++k) amax[k] = -DBL_MAX;
  bool first = true;
  for (int win = blockIdx.x; win < T.nwin; win += gridDim.x) {
    const int row = win * TILE_WIN_PART + tid;
    const bool valid = row < T.rows;
    const size_t o0 = (size_t)win * TILE_WIN_PART + tid;
    // only the blocks in which this window has entries (all of them on random sparsity; two or three
    // for the identity-like parts of a Lasso-type [P | A'])
    const int b0 = __ldg(T.wb_ptr + win), nbw = __ldg(T.wb_ptr + win + 1) - b0;
    const int ng = (nbw + TEPI_G - 1) / TEPI_G;
    const bool all = nbw == T.nblocks;
    double vn[TEPI_G]; unsigned short qn[TEPI_G];
    auto load = [&](int g) {
      if (all) {                                   // random sparsity: every block, no indirection
#pragma unroll
        for (int k = 0; k < TEPI_G; ++k) {
          const int c = g * TEPI_G + k;
          vn[k] = 0.0; qn[k] = 0;
          if (c < nbw) {
            const size_t o = (size_t)c * T.rows_pad + o0;
            vn[k] = __ldcg(partial + o); qn[k] = __ldg(T.inv + o);
          }
        }
      } else {
        int cb[TEPI_G];
#pragma unroll
        for (int k = 0; k < TEPI_G; ++k) cb[k] = (g * TEPI_G + k < nbw) ? __ldg(T.wb_blk + b0 + g * TEPI_G + k) : -1;
#pragma unroll
        for (int k = 0; k < TEPI_G; ++k) {
          vn[k] = 0.0; qn[k] = 0;
          if (cb[k] >= 0) {
            const size_t o = (size_t)cb[k] * T.rows_pad + o0;
            vn[k] = __ldcg(partial + o); qn[k] = __ldg(T.inv + o);
          }
        }
      }
    };
    if (ng > 0) load(0);
    typename Epi::Pre pre;
    if (valid) pre = epi.load(row);
    double sum = 0.0;
    for (int g = 0; g < ng; ++g) {
      if (!first) __syncthreads();                 // the previous group's picks are done
      first = false;
#pragma unroll
      for (int k = 0; k < TEPI_G; ++k) sp[k][tid] = vn[k];
      unsigned short q[TEPI_G];
#pragma unroll
      for (int k = 0; k < TEPI_G; ++k) q[k] = qn[k];
      if (g + 1 < ng) load(g + 1);
      __syncthreads();
#pragma unroll
      for (int k = 0; k < TEPI_G; ++k)
        if (g * TEPI_G + k < nbw) sum += sp[k][q[k]];
    }
    if (valid) epi.row(row, &sum, pre, asum, amax);
  }
  if (Epi::NSUM + Epi::NMAX > 0) {
    double *out = partials_out + (size_t)blockIdx.x * (Epi::NSUM + Epi::NMAX);
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
// CTAs of k_tile_epi = per-CTA partial rows it produces
static inline int tile_epi_grid(int nwin) { return nwin < TEPI_CTAS_PER_SM * OSQP_CUDA_SMS ? nwin : TEPI_CTAS_PER_SM * OSQP_CUDA_SMS; }

static inline TileView tile_view_of(const TileDev &D) {
  TileView V;
  V.rows = D.rows; V.rows_pad = D.rows_pad; V.cols = D.cols; V.cb = D.cb; V.nblocks = D.nblocks; V.nwin = D.nwin;
  V.win = D.win; V.ncta = D.ncta; V.racc = D.racc; V.nsb = D.nsb; V.n_lo = D.n_lo; V.cb_lo = D.cb_lo; V.nb_lo = D.nb_lo;
  V.slice_ptr = D.slice_ptr; V.inv = D.perm; V.perm = D.perm; V.col = D.col; V.val = D.val;
  V.item_ptr = D.item_ptr; V.item_blk = D.item_blk; V.item_ws = D.item_ws;
  V.cta_row = D.cta_row; V.cta_soff = D.cta_soff; V.wb_ptr = D.wb_ptr; V.wb_blk = D.wb_blk; V.ne = D.ne;
  return V;
}
static inline size_t tile_smem_bytes(const TileDev &D) {
  return (size_t)(D.nb_lo && D.cb_lo > D.cb ? D.cb_lo : D.cb) * 8 + (D.rows_mode ? (size_t)D.racc * 10 : 0);
}
// epilogues whose begin() may run twice per launch (tile kernel + epilogue kernel)
template <class E> struct epi_tile_ok { static constexpr bool value = false; };
