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
//  * val[] / col[] / rowptr[] are contiguous streams: one elected thread per CTA
//    copies tile t+2 into a three-deep shared-memory ring with TMA bulk copies
//    (cp.async.bulk + mbarrier complete_tx; SASS UBLKCP), which bypass the LSU /
//    L1TEX pipeline -- measured: mixing DRAM-latency streams and L2-hit gathers in
//    the L1TEX FIFO serialises them (profiles/r1_spmv_phase_ablation.log);
//  * every thread issues its gathers of the dense vector for tile t+1 back to back
//    (independent LDGs, L2 hits) BEFORE the row reduction of tile t, consumes them
//    after it, and parks the products in place of the values;
//  * groups of G lanes (G = 1..32 chosen from the mean row length: "vector per
//    row" / "warp per row") reduce each row segment out of shared memory in a
//    fixed order and run a fused row epilogue (rho scaling, ADMM updates,
//    residual norms, dot-product partials ...);
//  * rows longer than a tile are carried across tiles inside the CTA.
//
// NOT part of the product since round 2: the kernel below is the measured alternative of DESIGN.md
// section 6 (conclusion 3); the product includes csrc/spmv_common.cuh only.
#pragma once
#include "../../osqp_b200/csrc/spmv_common.cuh"

// One unit of work of a CTA: tile [ts, te) of chunk ck (rows [r0, r1), entries [s0, s1)).
struct SpmvItem {
  int ck, r0, r1, s0, s1, ts;
  bool valid;
  int4 dn;   // descriptor of chunk ck (before spmv_item_take) / ck + gridDim.x (after)
};
// Chunk descriptors are one 16-byte load each and are fetched one item ahead (dn): the
// two dependent global loads they replace sat on the critical path of every tile.
__device__ __forceinline__ void spmv_item_take(SpmvItem &it, const CsrView &A) {
  it.valid = false;
  while (it.ck < A.nchunks) {
    const int4 d = it.dn;
    const int nk = it.ck + gridDim.x;
    if (nk < A.nchunks) it.dn = __ldg(A.chunk_desc + nk);   // in flight until the next take
    if (d.x < d.y) {
      it.r0 = d.x; it.r1 = d.y; it.s0 = d.z; it.s1 = d.w;
      it.ts = it.s0;
      it.valid = true;
      return;
    }
    it.ck = nk;
  }
}
__device__ __forceinline__ void spmv_item_first(SpmvItem &it, const CsrView &A) {
  it.ck = blockIdx.x;
  it.dn = make_int4(0, 0, 0, 0);
  if (it.ck < A.nchunks) it.dn = __ldg(A.chunk_desc + it.ck);
  spmv_item_take(it, A);
}
__device__ __forceinline__ void spmv_item_next(SpmvItem &it, const CsrView &A) {
  if (it.ts + SPMV_TILE < it.s1) { it.ts += SPMV_TILE; return; }
  it.ck += gridDim.x;
  spmv_item_take(it, A);
}
// ---- TMA bulk copies (cp.async.bulk -> SASS UBLKCP) completing on an mbarrier.  The
// streams bypass the LSU / L1TEX pipeline, which is left to the gathers.
// One elected thread streams the tile of `it` into (scol, sval) and the row pointers of its
// chunk into srp (when they fit); all three copies complete on `bar`.
__device__ __forceinline__ void spmv_prefetch(const SpmvItem &it, const CsrView &A, int *scol,
                                              double *sval, int *srp, unsigned long long *bar) {
  if (it.valid && threadIdx.x == 0) {
    const int te = min(it.ts + SPMV_TILE, it.s1);
    const int a0 = it.ts & ~3;
    const unsigned cntp = (unsigned)((te - a0 + 3) & ~3);       // <= TILE + 4, arrays are padded
    const int rq0 = it.r0 & ~3;
    const unsigned nrpp = (unsigned)((it.r1 + 1 - rq0 + 3) & ~3);
    const bool staged = nrpp <= SPMV_RP_MAX;
    fence_proxy_async();          // generic-proxy accesses of this buffer precede the async writes
    mbar_expect_tx(bar, cntp * 12u + (staged ? nrpp * 4u : 0u));
    if (cntp) {
      bulk_g2s(scol, A.col + a0, cntp * 4u, bar);
      bulk_g2s(sval, A.val + a0, cntp * 8u, bar);
    }
    if (staged) bulk_g2s(srp, A.rowptr + rq0, nrpp * 4u, bar);
  }
}

// ---------------------------------------------------------------------------
// Warp-specialised kernel.  One CTA = 9 warps in three roles that only meet at
// shared-memory mbarriers, so the DRAM latency of the streams, the L2 latency of
// the gathers and the shuffle/epilogue latency of the row reduction overlap
// instead of adding up (profiles/r1_spmv_phase_ablation_*.log):
//   warp 8 (one lane)  PRODUCER  walks the chunk descriptors and TMA-copies tile
//                                t + STAGES - 1 into the ring       -> full[s]
//   warps 0-3          GATHER    col -> x[col] (L2-resident), products overwrite
//                                the values in place                -> ready[s]
//   warps 4-7          REDUCE    G lanes per row out of shared memory in a fixed
//                                order + fused row epilogue         -> free[s]
// ---------------------------------------------------------------------------
#define SPMV_STAGES         4
#define SPMV_GATHER_WARPS   4
#define SPMV_REDUCE_WARPS   4
#define SPMV_GATHER_THREADS (32 * SPMV_GATHER_WARPS)
#define SPMV_REDUCE_THREADS (32 * SPMV_REDUCE_WARPS)
#define SPMV_THREADS        (SPMV_GATHER_THREADS + SPMV_REDUCE_THREADS + 32)
#define SPMV_GATHER_SLOTS   ((SPMV_TILE + 4 + SPMV_GATHER_THREADS - 1) / SPMV_GATHER_THREADS)

// barrier among the reduce warps only (id 1; id 0 is __syncthreads)
__device__ __forceinline__ void reduce_warps_sync() {
  asm volatile("bar.sync 1, %0;\n" ::"n"(SPMV_REDUCE_THREADS) : "memory");
}

// per-stage item descriptor handed from the producer to the other roles
struct __align__(16) SpmvDesc {
  int ck, r0, r1, s0, s1, ts, valid, pad;
};

template <int NV, bool SPLIT, int MODE, class Epi>
__global__ void __launch_bounds__(SPMV_THREADS, (NV == 2 ? 2 : 3))
spmv_stream_kernel(CsrView A, Gather g0, Gather g1, Epi epi_in, double *partials) {
  constexpr int NS = SPLIT ? 2 : 1;
  constexpr int NACC = NV * NS;
  constexpr int ST = SPMV_STAGES;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double *sval0 = reinterpret_cast<double *>(smem_raw);                   // ST x TILE_PAD doubles
  double *prodB = sval0 + ST * SPMV_TILE_PAD;                             // NV == 2: ST x TILE_PAD
  double *carry = prodB + (NV == 2 ? ST * SPMV_TILE_PAD : 0);             // 2 * NACC
  double *scratch = carry + 2 * NACC;                                     // 32
  unsigned long long *bar_full = reinterpret_cast<unsigned long long *>(scratch + 32);   // ST
  unsigned long long *bar_ready = bar_full + ST;                          // ST
  unsigned long long *bar_free = bar_ready + ST;                          // ST
  SpmvDesc *sdesc = reinterpret_cast<SpmvDesc *>(bar_free + ST);          // ST (16-byte aligned)
  int *scol0 = reinterpret_cast<int *>(sdesc + ST);                       // ST x TILE_PAD ints
  int *srp0 = scol0 + ST * SPMV_TILE_PAD;                                 // ST x RP_MAX ints

  Epi epi = epi_in;
  if (!epi.begin(scratch)) return;   // same decision in every thread of every CTA
  const int dbg = g_spmv_dbg;
  const int tid = threadIdx.x;
  const int warp = tid >> 5;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < ST; ++s) {
      mbar_init(bar_full + s, 1);
      mbar_init(bar_ready + s, SPMV_GATHER_WARPS);
      mbar_init(bar_free + s, SPMV_REDUCE_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();

  double asum[Epi::NSUM > 0 ? Epi::NSUM : 1];
  double amax[Epi::NMAX > 0 ? Epi::NMAX : 1];
#pragma unroll
  for (int k = 0; k < Epi::NSUM; ++k) asum[k] = 0.0;
#pragma unroll
  for (int k = 0; k < Epi::NMAX; ++k) amax[k] = -DBL_MAX;

  if (warp == SPMV_GATHER_WARPS + SPMV_REDUCE_WARPS) {
    // =============================================================== PRODUCER
    if ((tid & 31) == 0) {
      SpmvItem item;
      spmv_item_first(item, A);
      for (int it = 0;; ++it) {
        const int s = it % ST;
        if (it >= ST) mbar_wait(bar_free + s, ((it / ST) - 1) & 1);   // reducers released the stage
        SpmvDesc d;
        d.ck = item.ck; d.r0 = item.r0; d.r1 = item.r1; d.s0 = item.s0; d.s1 = item.s1;
        d.ts = item.ts; d.valid = item.valid ? 1 : 0; d.pad = 0;
        sdesc[s] = d;
        if (!item.valid) { mbar_arrive(bar_full + s); break; }      // end marker, no copies
        const int te = min(item.ts + SPMV_TILE, item.s1);
        const int a0 = item.ts & ~3;
        const unsigned cntp = (unsigned)((te - a0 + 3) & ~3);       // <= TILE + 4, arrays are padded
        const int rq0 = item.r0 & ~3;
        const unsigned nrpp = (unsigned)((item.r1 + 1 - rq0 + 3) & ~3);
        const bool staged = nrpp <= SPMV_RP_MAX;
        fence_proxy_async();      // generic-proxy use of this stage precedes the async writes
        mbar_expect_tx(bar_full + s, cntp * 12u + (staged ? nrpp * 4u : 0u));
        if (cntp) {
          bulk_g2s(scol0 + s * SPMV_TILE_PAD, A.col + a0, cntp * 4u, bar_full + s);
          bulk_g2s(sval0 + s * SPMV_TILE_PAD, A.val + a0, cntp * 8u, bar_full + s);
        }
        if (staged) bulk_g2s(srp0 + s * SPMV_RP_MAX, A.rowptr + rq0, nrpp * 4u, bar_full + s);
        spmv_item_next(item, A);
      }
    }
  } else if (warp < SPMV_GATHER_WARPS) {
    // ================================================================= GATHER
    double x[SPMV_GATHER_SLOTS], y[NV == 2 ? SPMV_GATHER_SLOTS : 1];
    for (int it = 0;; ++it) {
      const int s = it % ST;
      mbar_wait(bar_full + s, (it / ST) & 1);
      const SpmvDesc d = sdesc[s];
      if (!d.valid) break;
      const int te = min(d.ts + SPMV_TILE, d.s1);
      const int a0 = d.ts & ~3;
      const int cnt = te - a0;
      const int *scol = scol0 + s * SPMV_TILE_PAD;
      double *prod = sval0 + s * SPMV_TILE_PAD;
      double *prod1 = prodB + s * SPMV_TILE_PAD;
      int cidx[SPMV_GATHER_SLOTS];
#pragma unroll
      for (int k = 0; k < SPMV_GATHER_SLOTS; ++k) {
        const int o = tid + k * SPMV_GATHER_THREADS;
        cidx[k] = (o < cnt) ? scol[o] : 0;
      }
#pragma unroll
      for (int k = 0; k < SPMV_GATHER_SLOTS; ++k) {
        const int o = tid + k * SPMV_GATHER_THREADS;
        x[k] = 0.0;
        if (NV == 2) y[k] = 0.0;
        if (o < cnt) {
          const int c = cidx[k];
          if (dbg & 1) { x[k] = 1.0; continue; }
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
#pragma unroll
      for (int k = 0; k < SPMV_GATHER_SLOTS; ++k) {
        const int o = tid + k * SPMV_GATHER_THREADS;
        if (o < cnt) {
          const double v = prod[o];
          if (MODE == 0) {
            prod[o] = v * x[k];
          } else {
            prod[o] = (cidx[k] < A.split_col) ? (x[k] < 0.0 ? v : 0.0) : v * v * x[k];
          }
          if (NV == 2) prod1[o] = v * y[k];
        }
      }
      __syncwarp();
      if ((tid & 31) == 0) mbar_arrive(bar_ready + s);
    }
  } else {
    // ================================================================= REDUCE
    const int rtid = tid - SPMV_GATHER_THREADS;
    const int G = A.group;
    const int lane_g = rtid & (G - 1);
    const int gid = rtid / G;
    const int ngroups = SPMV_REDUCE_THREADS / G;
    int ra = 0;          // first row of the chunk not yet finalised
    int par = 0;         // carry buffer parity
    int prev_ck = -1;
    for (int it = 0;; ++it) {
      const int s = it % ST;
      mbar_wait(bar_full + s, (it / ST) & 1);
      const SpmvDesc d = sdesc[s];
      if (!d.valid) break;
      mbar_wait(bar_ready + s, (it / ST) & 1);   // products of this tile are in place
      if (d.ck != prev_ck) { ra = d.r0; par = 0; prev_ck = d.ck; }
      const double *prod = sval0 + s * SPMV_TILE_PAD;
      const double *prod1 = prodB + s * SPMV_TILE_PAD;
      const int ts = d.ts, s1 = d.s1, r0 = d.r0, r1 = d.r1;
      const int rq0 = r0 & ~3;
      const bool staged = ((r1 + 1 - rq0 + 3) & ~3) <= SPMV_RP_MAX;
      const int *srp = srp0 + s * SPMV_RP_MAX;
      auto rowptr_at = [&](int r) { return staged ? srp[r - rq0] : A.rowptr[r]; };
      const int te = min(ts + SPMV_TILE, s1);
      const bool last = (te >= s1);
      const int a0 = ts & ~3;
      int rb = r1;
      if (!last) {
        // first row whose last entry lies at or beyond te (empty rows: rowptr >= te)
        int lo = ra, hi = r1;
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          const int b = rowptr_at(mid), e = rowptr_at(mid + 1);
          const bool fin = (e > b) ? (e <= te) : (b < te);
          if (fin) lo = mid + 1; else hi = mid;
        }
        rb = lo;
      }
      // the row rb is partially covered when this is not the last tile of the chunk
      const int rend = last ? rb : rb + 1;
      const int nrows = rend - ra;
      const int trips = (nrows + ngroups - 1) / ngroups;
      for (int t = 0; t < ((dbg & 2) ? 0 : trips); ++t) {
        const int r = ra + t * ngroups + gid;
        const bool valid = r < rend;
        const bool fin = valid && lane_g == 0 && r < rb;
        typename Epi::Pre pre;
        if (fin) pre = epi.load(r);          // latency overlaps the shared-memory reduction
        int lo = 0, hi = 0, sp = 0;
        if (valid) {
          lo = rowptr_at(r);
          hi = rowptr_at(r + 1);
          sp = SPLIT ? A.split[r] : hi;
        }
        const int clo = max(lo, ts), chi = min(hi, te);
        double sacc[NACC];
#pragma unroll
        for (int k = 0; k < NACC; ++k) sacc[k] = 0.0;
        for (int i = clo + lane_g; i < chi; i += G) {
          const double pv = prod[i - a0];
          if (SPLIT) {
            if (i < sp) sacc[0] += pv; else sacc[1] += pv;
          } else {
            sacc[0] += pv;
          }
          if (NV == 2) {
            const double pw = prod1[i - a0];
            if (SPLIT) {
              if (i < sp) sacc[NS] += pw; else sacc[NS + 1] += pw;
            } else {
              sacc[NS] += pw;
            }
          }
        }
#pragma unroll
        for (int k = 0; k < NACC; ++k)
          for (int o = G >> 1; o > 0; o >>= 1) sacc[k] += __shfl_xor_sync(0xffffffffu, sacc[k], o);
        if (valid && lane_g == 0) {
          if (r == ra && lo < ts) {
#pragma unroll
            for (int k = 0; k < NACC; ++k) sacc[k] += carry[par * NACC + k];
          }
          if (r < rb) {
            epi.row(r, sacc, pre, asum, amax);
          } else {
#pragma unroll
            for (int k = 0; k < NACC; ++k) carry[(par ^ 1) * NACC + k] = sacc[k];
          }
        }
      }
      if (!last) {           // CTA-uniform: a row continues in the next tile of this chunk
        reduce_warps_sync(); // carry written by one group, read by another
        ra = rb; par ^= 1;
      }
      __syncwarp();
      if ((tid & 31) == 0) mbar_arrive(bar_free + s);
    }
  }
  // ---- per-CTA partials of the fused reductions (only the reduce warps hold non-neutral values)
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
  return (size_t)((1 + (NV == 2 ? 1 : 0)) * SPMV_STAGES * SPMV_TILE_PAD + 2 * NV * (SPLIT ? 2 : 1) + 32) * sizeof(double) +
         (size_t)3 * SPMV_STAGES * sizeof(unsigned long long) + (size_t)SPMV_STAGES * sizeof(SpmvDesc) +
         (size_t)SPMV_STAGES * (SPMV_TILE_PAD + SPMV_RP_MAX) * sizeof(int);
}

