// spmv_sell.cuh -- sliced-ELL mirror of a CSR matrix and the SpMV kernel that streams it.
//
// Replaces, like spmv_direct.cuh, the CSC loops of reference src/lin_alg.c:241-322
// (mat_vec / mat_tpose_vec) and, through the row epilogues, the vector loops of
// src/auxil.c:161-225.  Same template signature and epilogue concept as the other two kernels.
//
// Why a third layout: the direct CSR kernel runs AT the L1TEX sector rate (~1 sector/clk/SM,
// DESIGN section 6): per 32 nonzeros it pays 32 gather sectors plus ~21 sectors for col/val
// (row segments of 10 entries read by 8-lane groups straddle sectors) plus row pointers and
// shuffles.  Here
//   * the nonzeros are stored pass-major: a SLICE is 32/G consecutive rows (G lanes per row,
//     G = 1 for short rows), and pass p of a slice is ONE aligned run of 32 col / 32 val
//     entries, lane l owning entry p*G + (l % G) of row l / G.  Rows of a slice are padded to
//     the longest one (col = -1, val = 0).  With G = 1 there is no cross-lane reduction and
//     no row pointer at all; lane l IS row l of the slice, so the fused epilogue is coalesced.
//     G is chosen per BLOCK of 32 rows (the value that pads the block least), so matrices
//     that mix 500-entry and 2-entry rows (Lasso, identity blocks) neither pad the short rows
//     to a warp nor walk the long ones with one lane;
//   * every warp of a persistent grid owns a CONTIGUOUS range of slices (balanced by passes,
//     table built at setup), i.e. one contiguous run of the col stream and one of the val
//     stream, read in chunks of CH passes with fully coalesced evict-first loads: 12 sectors
//     per 32 nonzeros, the minimum.  All CH gathers of a lane are issued back to back;
//   * no CTA-level synchronisation and no shared memory on the row path.
// Measured (tools/microbench/spmv_sell_bench.cu, profiles/r1_sell_microbench.log): 43 us
// against 55 us for the direct CSR kernel on config 2's A, 50 against 60 us on its M.
// SELL_TMA=1 keeps the alternative that was tried first: lane 0 copies each chunk into a
// warp-private shared-memory ring with TMA bulk copies (cp.async.bulk + mbarrier
// complete_tx, SASS UBLKCP) and the gathers of chunk c+1 are issued before chunk c is
// consumed.  It removes the col/val sectors from L1TEX but is never faster than 47 us, and
// it degrades sharply as the rings grow (86 us at 216 KB of shared memory per SM): what the
// gathers need is L1 capacity for lines in flight, and the carve-out takes it away.
// Sums depend on (G, slice order) only -> bit-reproducible run to run.
//
// The mirror is built from the CSR arrays on the device (sell_build) and its values are
// re-read from them after every scaling / update (sell_refresh); the CSR copy stays the
// canonical one for setup, scaling, matrix updates and the small-problem fused kernel.
#pragma once
#include "spmv_common.cuh"

#ifndef SELL_CH
#define SELL_CH 4          // passes (of 32 entries) per chunk = independent gathers per lane in flight
#endif
#ifndef SELL_NST
#define SELL_NST 3         // chunks in flight per warp
#endif
#ifndef SELL_MINB
#define SELL_MINB 4        // resident CTAs per SM (8 warps each; 64 registers per thread)
#endif
#ifndef SELL_TMA
#define SELL_TMA 0         // 1: stream col/val through warp-private TMA rings (ablation, slower)
#endif
#ifndef SELL_PF
#define SELL_PF 1          // prefetch the next chunk's columns behind the current gathers
#endif
#ifndef SELL_STREAM_LD
#define SELL_STREAM_LD 0   // col/val loads: 0 evict-first (ld.global.cs), 1 L1::no_allocate
#endif
__device__ __forceinline__ int sell_ld_col(const int *p) {
#if SELL_STREAM_LD == 1
  int v; asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p)); return v;
#else
  return __ldcs(p);
#endif
}
__device__ __forceinline__ double sell_ld_val(const double *p) {
#if SELL_STREAM_LD == 1
  double v; asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p)); return v;
#else
  return __ldcs(p);
#endif
}
#define SELL_WARPS (OSQP_CUDA_THREADS / 32)
#define SELL_STAGE_BYTES (SELL_CH * 32 * 12)

struct SellView {
  int rows, nslices, split_col;
  const int *slice_ptr;     // nslices + 1, in passes
  const int *slice_info;    // nslices: (first row << 3) | log2(lanes per row)
  const int *warp_slice;    // (grid * SELL_WARPS) + 1: first slice of every warp
  const int *col;           // total_passes * 32
  const double *val;
};


// One nonzero into the row accumulators; shared by spmv_sell_kernel and spmv_sell_rows so that
// both round identically (explicit fma, same operand order).  Padding arrives as v = 0, x = 0.
template <int NV, bool SPLIT, int MODE>
__device__ __forceinline__ void sell_accumulate(double *acc, double v, double xj, double yj, int cc,
                                                int row, int split_col) {
  constexpr int NS = SPLIT ? 2 : 1;
  const bool hi = SPLIT && cc >= split_col;
  if (MODE == 0) {
    if (hi) acc[NS - 1] = fma(v, xj, acc[NS - 1]); else acc[0] = fma(v, xj, acc[0]);
  } else {
    const double pv = (cc < split_col) ? ((cc == row) ? v : 0.0) : v * v * xj;
    if (hi) acc[NS - 1] += pv; else acc[0] += pv;
  }
  if (NV == 2) {
    if (hi) acc[2 * NS - 1] = fma(v, yj, acc[2 * NS - 1]); else acc[NS] = fma(v, yj, acc[NS]);
  }
}

template <int NV, bool SPLIT, int MODE, class Epi>
__global__ void __launch_bounds__(OSQP_CUDA_THREADS, (NV == 2 ? 2 : SELL_MINB))
spmv_sell_kernel(SellView A, Gather g0, Gather g1, Epi epi_in, double *partials) {
  constexpr int NS = SPLIT ? 2 : 1;
  constexpr int NACC = NV * NS;
  constexpr int CH = SELL_CH, NST = SELL_NST;
#if SELL_TMA
  extern __shared__ __align__(128) unsigned char sell_smem[];
#endif
  __shared__ double scratch[32];
#if SELL_TMA
  __shared__ unsigned long long bars[SELL_WARPS * NST];
#endif
  Epi epi = epi_in;
  if (!epi.begin(scratch)) return;   // same decision in every thread of every CTA
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  double asum[Epi::NSUM > 0 ? Epi::NSUM : 1];
  double amax[Epi::NMAX > 0 ? Epi::NMAX : 1];
#pragma unroll
  for (int k = 0; k < Epi::NSUM; ++k) asum[k] = 0.0;
#pragma unroll
  for (int k = 0; k < Epi::NMAX; ++k) amax[k] = -DBL_MAX;

  const int gw = blockIdx.x * SELL_WARPS + warp;
  int s = __ldg(A.warp_slice + gw);
  const int s_end = __ldg(A.warp_slice + gw + 1);
  const int P0 = __ldg(A.slice_ptr + s);
  const int P1 = __ldg(A.slice_ptr + s_end);
  const int nch = (P1 - P0 + CH - 1) / CH;
#if SELL_TMA
  unsigned char *wbase = sell_smem + (size_t)warp * (NST * SELL_STAGE_BYTES);
  unsigned long long *wbar = bars + warp * NST;
#endif

#if SELL_TMA
  auto issue = [&](int ch) {   // lane 0 only
    const int st = ch % NST;
    const int p = P0 + ch * CH;
    const unsigned np = (unsigned)min(CH, P1 - p);
    unsigned char *sb = wbase + st * SELL_STAGE_BYTES;
    fence_proxy_async();       // generic-proxy reads of this stage precede the async writes
    mbar_expect_tx(wbar + st, np * 384u);
    bulk_g2s(sb, A.col + (size_t)p * 32, np * 128u, wbar + st);
    bulk_g2s(sb + CH * 128, A.val + (size_t)p * 32, np * 256u, wbar + st);
  };
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < NST; ++k) mbar_init(wbar + k, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    for (int k = 0; k < NST && k < nch; ++k) issue(k);
  }
  __syncwarp();
#endif

  // ---- slice state.  p1 = slice_ptr[s + 1], p2 = slice_ptr[s + 2] (fetched one slice ahead)
  double acc[NACC];
#pragma unroll
  for (int k = 0; k < NACC; ++k) acc[k] = 0.0;
  int pc = P0;
  int p1 = (s < s_end) ? __ldg(A.slice_ptr + s + 1) : P0;
  int p2 = (s + 1 < s_end) ? __ldg(A.slice_ptr + s + 2) : p1;
  int left = p1 - pc;
  // i1 = slice_info[s], i2 = slice_info[s + 1] (fetched one slice ahead, like p2)
  int i1 = (s < s_end) ? __ldg(A.slice_info + s) : 0;
  int i2 = (s + 1 < s_end) ? __ldg(A.slice_info + s + 1) : 0;
  int lg = i1 & 7;
  int row = (i1 >> 3) + (lane >> lg);
  bool mine = (s < s_end) && ((lane & ((1 << lg) - 1)) == 0) && (row < A.rows);
  typename Epi::Pre pre;
  if (mine) pre = epi.load(row);

  auto finish = [&]() {        // close slice s, open slice s + 1
    for (int o = (1 << lg) >> 1; o > 0; o >>= 1) {
#pragma unroll
      for (int k = 0; k < NACC; ++k) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
    }
    if (mine) epi.row(row, acc, pre, asum, amax);
#pragma unroll
    for (int k = 0; k < NACC; ++k) acc[k] = 0.0;
    ++s;
    pc = p1; p1 = p2; i1 = i2;
    if (s + 1 < s_end) { p2 = __ldg(A.slice_ptr + s + 2); i2 = __ldg(A.slice_info + s + 1); }
    left = p1 - pc;
    lg = i1 & 7;
    row = (i1 >> 3) + (lane >> lg);
    mine = (s < s_end) && ((lane & ((1 << lg) - 1)) == 0) && (row < A.rows);
    if (mine) pre = epi.load(row);
  };
  while (s < s_end && left == 0) finish();   // leading slices of empty rows

  // one accumulation step of pass j (value v, gathered x / y, column cc) + slice bookkeeping
  auto accumulate = [&](double v, double xj, double yj, int cc) {
    sell_accumulate<NV, SPLIT, MODE>(acc, v, xj, yj, cc, row, A.split_col);
    if (--left == 0) { do { finish(); } while (s < s_end && left == 0); }
  };
  // gathers of one entry; padding (cc < 0, val = 0) gathers nothing
  auto gather1 = [&](int cc, double &xj, double &yj) {
    xj = 0.0; yj = 0.0;
    if (cc >= 0) {
      if (MODE == 0) {
        xj = ldg_issue((cc < A.split_col) ? g0.lo + cc : g0.hi + (cc - A.split_col));
        if (NV == 2) yj = ldg_issue((cc < A.split_col) ? g1.lo + cc : g1.hi + (cc - A.split_col));
      } else if (cc >= A.split_col) {
        xj = ldg_issue(g0.hi + (cc - A.split_col));   // Jacobi: val^2 * rho on the A' part
      }
    }
  };
  constexpr int NY = (NV == 2) ? CH : 1;
  constexpr bool NEED_COL = SPLIT || MODE != 0;

#if SELL_TMA
  // Software pipeline inside the warp: the gathers of chunk ch + 1 are issued (columns read
  // from its stage) BEFORE chunk ch is consumed (values read from its stage, stage refilled),
  // so that a warp always has a chunk of L2-latency gathers in flight.
  double xa[CH], xb[CH], ya[NY], yb[NY];
  auto gather = [&](int ch, double (&x)[CH], double (&y)[NY]) {
    if (ch >= nch) return;
    const int st = ch % NST;
    mbar_wait(wbar + st, (ch / NST) & 1);
    const int *scol = reinterpret_cast<const int *>(wbase + st * SELL_STAGE_BYTES);
    const int np = min(CH, P1 - P0 - ch * CH);
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      const int cc = (j < np) ? scol[j * 32 + lane] : -1;
      double yj;
      gather1(cc, x[j], yj);
      if (NV == 2) y[j] = yj;
    }
  };
  auto consume = [&](int ch, double (&x)[CH], double (&y)[NY]) {
    const int st = ch % NST;
    const int *scol = reinterpret_cast<const int *>(wbase + st * SELL_STAGE_BYTES);
    const double *sval = reinterpret_cast<const double *>(wbase + st * SELL_STAGE_BYTES + CH * 128);
    const int np = min(CH, P1 - P0 - ch * CH);
    double v[CH];
    int c[NEED_COL ? CH : 1];
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      v[j] = (j < np) ? sval[j * 32 + lane] : 0.0;
      if (NEED_COL) c[j] = (j < np) ? scol[j * 32 + lane] : -1;
    }
    __syncwarp();              // every lane has read the stage: refill it
    if (lane == 0 && ch + NST < nch) issue(ch + NST);
#pragma unroll
    for (int j = 0; j < CH; ++j)
      if (j < np) accumulate(v[j], x[j], (NV == 2) ? y[j] : 0.0, NEED_COL ? c[j] : 0);
  };
  gather(0, xa, ya);
  for (int ch = 0; ch < nch; ch += 2) {
    gather(ch + 1, xb, yb);
    consume(ch, xa, ya);
    gather(ch + 2, xa, ya);
    if (ch + 1 < nch) consume(ch + 1, xb, yb);
  }
#else
  // The columns of chunk ch + 1 are fetched while the gathers of chunk ch are in flight, so
  // that per chunk a warp waits for max(HBM stream latency, L2 gather latency), not the sum.
  int cn[CH];
  auto load_cols = [&](int ch) {
    const int np = min(CH, P1 - P0 - ch * CH);
    const size_t e0 = ((size_t)(P0 + ch * CH)) * 32 + lane;
#pragma unroll
    for (int j = 0; j < CH; ++j) cn[j] = (j < np) ? sell_ld_col(A.col + e0 + j * 32) : -1;
  };
  if (nch > 0) load_cols(0);
  for (int ch = 0; ch < nch; ++ch) {
    const int np = min(CH, P1 - P0 - ch * CH);
    int c[CH];
    double v[CH], x[CH], y[NY];
    const size_t e0 = ((size_t)(P0 + ch * CH)) * 32 + lane;
#pragma unroll
    for (int j = 0; j < CH; ++j) c[j] = cn[j];
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      double yj;
      gather1(c[j], x[j], yj);
      if (NV == 2) y[j] = yj;
    }
#pragma unroll
    for (int j = 0; j < CH; ++j) v[j] = (j < np) ? sell_ld_val(A.val + e0 + j * 32) : 0.0;
#if SELL_PF
    if (ch + 1 < nch) load_cols(ch + 1);
#endif
#pragma unroll
    for (int j = 0; j < CH; ++j)
      if (j < np) accumulate(v[j], x[j], (NV == 2) ? y[j] : 0.0, c[j]);
#if !SELL_PF
    if (ch + 1 < nch) load_cols(ch + 1);
#endif
  }
#endif

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
static inline SellView sell_view_of(const CsrDev &C) {
  SellView V;
  V.rows = C.rows; V.nslices = C.sell.ok ? C.sell.nslices : 0; V.split_col = C.split_col;
  V.slice_ptr = C.sell.slice_ptr; V.slice_info = C.sell.slice_info; V.warp_slice = C.sell.warp_slice;
  V.col = C.sell.col; V.val = C.sell.val;
  return V;
}
static inline size_t sell_smem_bytes() {
  return SELL_TMA ? (size_t)SELL_WARPS * SELL_NST * SELL_STAGE_BYTES : 0;
}

// ---------------------------------------------------------------------------------------------
// Warp-level body for kernels that own a whole (small) problem with one thread block
// (admm_fused.cuh, resid_batch_kernel): slices first_slice, first_slice + slice_stride, ... are
// walked by the calling warp; the gathered vectors may live in shared memory (generic loads).
// Same per-row summation order as spmv_sell_kernel (passes in order, then the xor tree over the
// G lanes), so a row sum is the same bits whichever of the two computes it.  The direct CSR
// body spends ~5 warp instructions per nonzero on 20-entry rows (index arithmetic, predicates
// and shuffles: ncu on the MPC problems of BASELINE config 4, DFMA = 1.2 % of the issue slots);
// here a pass of 32 nonzeros is two coalesced loads, one gather and one DFMA per lane.
// SMEM = true: the gathered vectors are in SHARED memory (admm_fused.cuh).  The gathers are then
// explicit ld.shared with 32-bit addresses, branch-free (padding reads element 0 and is
// discarded): the generic-pointer form cost ~22 instructions per gather (address-space checks,
// a divergent branch per select), more than half of the fused kernel's issue slots.
template <int NV, bool SPLIT, int MODE, bool SMEM = false, class Epi>
__device__ __forceinline__ void spmv_sell_rows(const SellView &A, const Gather &g0, const Gather &g1,
                                               Epi &epi, double *asum, double *amax,
                                               int first_slice, int slice_stride) {
  constexpr int NS = SPLIT ? 2 : 1;
  constexpr int NACC = NV * NS;
  constexpr int U = 4;
  const int lane = threadIdx.x & 31;
  unsigned s0lo = 0, s0hi = 0, s1lo = 0, s1hi = 0;
  if (SMEM) {
    s0lo = smem_u32(g0.lo); s0hi = g0.hi ? smem_u32(g0.hi) : 0u;
    if (NV == 2) { s1lo = smem_u32(g1.lo); s1hi = g1.hi ? smem_u32(g1.hi) : 0u; }
  }
  (void)s0lo; (void)s0hi; (void)s1lo; (void)s1hi;
  int s = first_slice;
  // metadata of the next slice is fetched while the current one is processed
  int np0 = 0, np1 = 0, ninfo = 0;
  if (s < A.nslices) { np0 = __ldg(A.slice_ptr + s); np1 = __ldg(A.slice_ptr + s + 1); ninfo = __ldg(A.slice_info + s); }
  while (s < A.nslices) {
    const int p0 = np0, p1 = np1, info = ninfo;
    const int sn = s + slice_stride;
    if (sn < A.nslices) { np0 = __ldg(A.slice_ptr + sn); np1 = __ldg(A.slice_ptr + sn + 1); ninfo = __ldg(A.slice_info + sn); }
    const int lg = info & 7;
    const int row = (info >> 3) + (lane >> lg);
    const bool mine = ((lane & ((1 << lg) - 1)) == 0) && (row < A.rows);
    double acc[NACC];
#pragma unroll
    for (int k = 0; k < NACC; ++k) acc[k] = 0.0;
    for (int p = p0; p < p1; p += U) {
      int c[U];
      double v[U];
#pragma unroll
      for (int j = 0; j < U; ++j) {
        const bool in = p + j < p1;
        c[j] = in ? __ldg(A.col + (size_t)(p + j) * 32 + lane) : -1;
        v[j] = in ? __ldg(A.val + (size_t)(p + j) * 32 + lane) : 0.0;
      }
      double x[U], y[NV == 2 ? U : 1];
      if (SMEM) {
        static_assert(!SMEM || MODE == 0, "shared-memory gathers: plain products only");
#pragma unroll
        for (int j = 0; j < U; ++j) {
          const int cc = c[j], ci = max(cc, 0);
          const bool lo = ci < A.split_col;
          const unsigned off = 8u * (unsigned)(lo ? ci : ci - A.split_col);
          double t;
          asm volatile("ld.shared.f64 %0, [%1];" : "=d"(t) : "r"((lo ? s0lo : s0hi) + off) : "memory");
          x[j] = cc >= 0 ? t : 0.0;
          if (NV == 2) {
            asm volatile("ld.shared.f64 %0, [%1];" : "=d"(t) : "r"((lo ? s1lo : s1hi) + off) : "memory");
            y[j] = cc >= 0 ? t : 0.0;
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < U; ++j) {
          const int cc = c[j];
          x[j] = 0.0;
          if (NV == 2) y[j] = 0.0;
          if (cc >= 0) {
            if (MODE == 0) {
              x[j] = *((cc < A.split_col) ? g0.lo + cc : g0.hi + (cc - A.split_col));
              if (NV == 2) y[j] = *((cc < A.split_col) ? g1.lo + cc : g1.hi + (cc - A.split_col));
            } else if (cc >= A.split_col) {
              x[j] = *(g0.hi + (cc - A.split_col));
            }
          }
        }
      }
      // entries beyond p1 arrive as v = 0, x = 0: fma(0, 0, acc) = acc, no predicate needed
#pragma unroll
      for (int j = 0; j < U; ++j)
        sell_accumulate<NV, SPLIT, MODE>(acc, v[j], x[j], (NV == 2) ? y[j] : 0.0, c[j], row, A.split_col);
    }
    for (int o = (1 << lg) >> 1; o > 0; o >>= 1) {
#pragma unroll
      for (int k = 0; k < NACC; ++k) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
    }
    if (mine) {
      typename Epi::Pre pre = epi.load(row);
      epi.row(row, acc, pre, asum, amax);
    }
    s = sn;
  }
}
