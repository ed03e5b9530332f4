// spmv_direct.cuh -- row-wise CSR SpMV for B200, "G lanes per row, straight from L2/HBM".
//
// Replaces the CSC scatter/gather loops of reference src/lin_alg.c:241-322
// (mat_vec / mat_tpose_vec) and, through the row epilogues, the vector loops of
// src/auxil.c:161-225 that consume their results.  Same template signature and
// epilogue concept as the mirrors (spmv_common.cuh); engine.cu (launch_stream) picks the kernel.
//
// Why a second kernel: on B200 the SpMV of a random-sparsity QP is bound by the L1TEX
// gather rate (one 8-byte gather of the dense vector per nonzero; measured ceiling
// ~1 gather/clk/SM, tools/microbench/gather_rate.cu), not by HBM.  Every extra L1TEX
// wavefront (shared-memory staging of col/val, parking the products, reading them back,
// shuffles) is paid on the same pipe -- the phase ablation in profiles/ shows the costs
// ADD instead of overlapping.  This kernel therefore touches L1TEX as little as possible:
//   * groups of G lanes (G = 1..32 from the mean row length) own TB rows at a time and walk
//     them in lockstep passes of G entries: each lane reads col/val with plain coalesced
//     loads, gathers the dense vector and accumulates in registers -- no shared memory at
//     all, and TB independent loads per lane in flight at every link of the dependent
//     chain rowptr -> col -> x[col] (the kernel is latency-bound otherwise);
//   * one xor-shuffle tree per row, then a shuffle "transpose" so that lane j of the warp
//     holds the sum of row (block base + j): the fused row epilogue (rho scaling, ADMM
//     updates, residual norms, dot-product partials ...) runs ONCE per block of rows with
//     fully coalesced vector loads/stores, and its inputs are prefetched (Epi::load) before
//     the gathers are issued;
//   * blocks of rows are dealt round-robin to the warps of a persistent grid (a multiple of
//     the 148 SMs); the summation order inside a row depends on G only -> bit-reproducible.
#pragma once
#include "spmv_common.cuh"

#ifndef SPMV_DIRECT_TB
#define SPMV_DIRECT_TB 4   // rows per lane group in flight: independent loads per lane at each step
#endif
#ifndef SPMV_DIRECT_MINB
#define SPMV_DIRECT_MINB 4 // resident CTAs per SM the register budget is set for
#endif

// col / val are read exactly once per launch: with SPMV_DIRECT_STREAM_CS they are loaded
// evict-first (ld.global.cs) so that they do not push the gathered vector out of L1 / L2
#ifndef SPMV_DIRECT_STREAM_CS
#define SPMV_DIRECT_STREAM_CS 1
#endif
template <class T> __device__ __forceinline__ T ld_stream(const T *p) {
#if SPMV_DIRECT_STREAM_CS
  return __ldcs(p);
#else
  return __ldg(p);
#endif
}

// gather of the dense vector: read-only global path, or a generic load when the vector
// lives in shared memory (admm_fused.cuh)
template <bool GEN> __device__ __forceinline__ double ld_gather(const double *p) {
  if (GEN) return *p;
  return __ldg(p);
}

// Row blocks first_block, first_block + block_stride, ... of A, processed by the calling WARP
// (no CTA-level synchronisation inside).  GEN: the gathered vectors may be in shared memory.
template <int NV, bool SPLIT, int MODE, bool GEN, class Epi>
__device__ __forceinline__ void spmv_direct_rows(const CsrView &A, const Gather &g0, const Gather &g1,
                                                 Epi &epi, double *asum, double *amax,
                                                 int first_block, int block_stride) {
  constexpr int NS = SPLIT ? 2 : 1;
  constexpr int NACC = NV * NS;
  constexpr int TB = SPMV_DIRECT_TB;
  const int lane = threadIdx.x & 31;
  const int G = A.group;                 // lanes per row (power of two)
  const int lg = 31 - __clz(G);
  const int RPT = 32 >> lg;              // rows per trip of the warp
  const int lane_g = lane & (G - 1);
  const int grp = lane >> lg;
  const int RB = (RPT >= 4) ? 32 : 8 * RPT;   // rows per block (one epilogue call per block)
  const int T = RB / RPT;                // trips per block
  const int src_lane = (lane & (RPT - 1)) << lg;   // a lane of the group that owns "my" row
  const int my_trip = lane >> (5 - lg);             // lane / RPT
  const int nblocks = (A.rows + RB - 1) / RB;

  for (int b = first_block; b < nblocks; b += block_stride) {
    const int base = b * RB;
    const int myrow = base + lane;
    const bool mine_valid = lane < RB && myrow < A.rows;
    typename Epi::Pre pre;
    if (mine_valid) pre = epi.load(myrow);   // latency overlaps the gathers below
    double mine[NACC];
#pragma unroll
    for (int k = 0; k < NACC; ++k) mine[k] = 0.0;
    const int rown = min(RB, A.rows - base);             // rows of this block
    const int rp = (lane <= rown) ? __ldg(A.rowptr + base + lane) : 0;   // one coalesced load
    const int rp_hi = (rown == 32) ? __ldg(A.rowptr + base + 32) : 0;    // 33rd pointer (uniform)

    // TB rows per lane group advance in lockstep passes of G entries: TB independent
    // rowptr / col / val / gather loads are in flight per lane at every step of the chain.
    for (int tb = 0; tb < T; tb += TB) {
      int lo[TB], hi[TB], sp[SPLIT ? TB : 1];
      int len_max = 0;
#pragma unroll
      for (int q = 0; q < TB; ++q) {
        const int ri = (tb + q) * RPT + grp;            // row of this lane group inside the block
        const bool ok = (tb + q < T) && (ri < rown);
        // row pointers come from the block's coalesced load (rp / rp_hi), not from 2 loads per row
        const int a = __shfl_sync(0xffffffffu, rp, ri & 31);
        const int b = __shfl_sync(0xffffffffu, rp, (ri + 1) & 31);
        lo[q] = ok ? a : 0;
        hi[q] = ok ? ((ri + 1 < 32) ? b : rp_hi) : 0;
        if (SPLIT) sp[q] = ok ? __ldg(A.split + base + ri) : 0;
      }
      double s[TB][NACC];
#pragma unroll
      for (int q = 0; q < TB; ++q) {
        len_max = max(len_max, hi[q] - lo[q]);
#pragma unroll
        for (int k = 0; k < NACC; ++k) s[q][k] = 0.0;
      }
      const int passes = ((int)__reduce_max_sync(0xffffffffu, (unsigned)len_max) + G - 1) >> lg;
      for (int p = 0; p < passes; ++p) {
        int c[TB];
        double v[TB], x[TB], y[NV == 2 ? TB : 1];
#pragma unroll
        for (int q = 0; q < TB; ++q) {
          const int i = lo[q] + lane_g + (p << lg);
          c[q] = (i < hi[q]) ? ld_stream(A.col + i) : -1;
        }
#pragma unroll
        for (int q = 0; q < TB; ++q) {
          const int i = lo[q] + lane_g + (p << lg);
          v[q] = 0.0; x[q] = 0.0;
          if (NV == 2) y[q] = 0.0;
          if (c[q] >= 0) {
            v[q] = ld_stream(A.val + i);
            const int cc = c[q];
            if (MODE == 0) {
              x[q] = ld_gather<GEN>((cc < A.split_col) ? g0.lo + cc : g0.hi + (cc - A.split_col));
            } else {
              // Jacobi: the P part keeps its diagonal entry, the A' part val^2 * rho
              const int r = base + (tb + q) * RPT + grp;
              if (cc < A.split_col) { x[q] = (cc == r) ? 1.0 : 0.0; }
              else { x[q] = v[q] * ld_gather<GEN>(g0.hi + (cc - A.split_col)); }
            }
            if (NV == 2) y[q] = ld_gather<GEN>((cc < A.split_col) ? g1.lo + cc : g1.hi + (cc - A.split_col));
          }
        }
#pragma unroll
        for (int q = 0; q < TB; ++q) {
          const int i = lo[q] + lane_g + (p << lg);
          if (c[q] >= 0) {
            const double pv = v[q] * x[q];
            if (SPLIT) {
              if (i < sp[q]) s[q][0] += pv; else s[q][1] += pv;
            } else {
              s[q][0] += pv;
            }
            if (NV == 2) {
              const double pw = v[q] * y[q];
              if (SPLIT) {
                if (i < sp[q]) s[q][NS] += pw; else s[q][NS + 1] += pw;
              } else {
                s[q][NS] += pw;
              }
            }
          }
        }
      }
      // fixed-order tree over the G lanes of each row, then hand the sum to the lane that
      // runs the epilogue of that row
      for (int o = G >> 1; o > 0; o >>= 1) {
#pragma unroll
        for (int q = 0; q < TB; ++q) {
#pragma unroll
          for (int k = 0; k < NACC; ++k) s[q][k] += __shfl_xor_sync(0xffffffffu, s[q][k], o);
        }
      }
#pragma unroll
      for (int q = 0; q < TB; ++q) {
#pragma unroll
        for (int k = 0; k < NACC; ++k) {
          const double tr = __shfl_sync(0xffffffffu, s[q][k], src_lane);
          if (my_trip == tb + q) mine[k] = tr;
        }
      }
    }
    if (mine_valid) epi.row(myrow, mine, pre, asum, amax);
  }
}

template <int NV, bool SPLIT, int MODE, class Epi>
__global__ void __launch_bounds__(OSQP_CUDA_THREADS, (NV == 2 ? 2 : SPMV_DIRECT_MINB))
spmv_direct_kernel(CsrView A, Gather g0, Gather g1, Epi epi_in, double *partials) {
  __shared__ double scratch[32];
  Epi epi = epi_in;
  if (!epi.begin(scratch)) return;   // same decision in every thread of every CTA
  const int tid = threadIdx.x;
  double asum[Epi::NSUM > 0 ? Epi::NSUM : 1];
  double amax[Epi::NMAX > 0 ? Epi::NMAX : 1];
#pragma unroll
  for (int k = 0; k < Epi::NSUM; ++k) asum[k] = 0.0;
#pragma unroll
  for (int k = 0; k < Epi::NMAX; ++k) amax[k] = -DBL_MAX;
  spmv_direct_rows<NV, SPLIT, MODE, false>(A, g0, g1, epi, asum, amax,
                                           blockIdx.x * (OSQP_CUDA_THREADS / 32) + (tid >> 5),
                                           gridDim.x * (OSQP_CUDA_THREADS / 32));

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

// lanes per row for the direct kernel: the smallest power of two that covers the mean row
// length in at most two passes of one load each (U loads are in flight anyway)
static inline int spmv_direct_group(int nnz, int rows) {
  const double avg = rows > 0 ? (double)nnz / rows : 0.0;
  int g = 1;
  while (g < 32 && avg >= 1.5 * g) g <<= 1;
  const char *e = getenv("OSQP_B200_GMUL");   // experiment knob
  if (e) { g *= atoi(e); if (g > 32) g = 32; if (g < 1) g = 1; }
  return g;
}
