// admm_fused.cuh -- small problems: `count` whole ADMM iterations (rhs, the complete CG loop,
// x / z / y updates) in ONE kernel launch by ONE thread block, with every iterate and CG
// vector in shared memory (5 n + 5 m doubles <= 220 KB).  Same mathematics as the
// multi-kernel path of engine.cu (reference src/osqp.c:354-370, src/auxil.c:161-225 with the
// KKT solve replaced by the Jacobi-PCG on P + sigma I + A' rho A); what changes is who
// synchronises: __syncthreads instead of kernel boundaries, CTA-wide reductions instead of
// per-CTA partials, and an exact device-side CG loop instead of enqueued launches that turn
// into no-ops.  A problem of BASELINE config 4 (MPC, n = 1520, m = 2540) costs ~70 launches
// of a few microseconds per ADMM iteration on the multi-kernel path and keeps 1 SM busy here,
// so osqp_solve_batch can run one problem per SM.
#pragma once
#include "spmv_direct.cuh"
#include "pcg.cuh"
#include "band.cuh"

#define ADMM_FUSED_THREADS 1024
#define ADMM_FUSED_MAX_SMEM (220 * 1024)

struct FusedArgs {
  CsrView A, M;
  SellView As, Ms;      // sliced-ELL mirrors (spmv_sell.cuh); use_sell = 0: direct CSR rows
  int use_sell;
  int n, m;
  double *x, *z, *y, *Ax, *v, *dx, *dy;          // iterates in global memory (in / out)
  const double *q, *l, *u, *rho, *Minv;
  PcgScal *S;
  double sigma, alpha, rel, absf;
  int maxit, use_cap, count, store;
  // exact block-tridiagonal preconditioner (band.cuh) instead of Jacobi, when the problem qualifies
  int use_band;
  BandView band;
  unsigned smem_bytes;  // dynamic shared memory this problem needs (iterates + band scratch)
};

static inline size_t admm_fused_smem(int n, int m) { return sizeof(double) * (5 * (size_t)n + 5 * (size_t)m); }

// ---- row epilogues on shared-memory vectors
struct FEpiRhs {     // r = -(q + [P|A'][x; v]), p = Minv r, xt = x
  static constexpr int NSUM = 1, NMAX = 1;
  const double *q, *Minv, *x;
  double *r, *p, *xt;
  EPI_NO_PREFETCH
  __device__ __forceinline__ void row(int j, const double *s, const Pre &, double *asum, double *amax) const {
    const double rj = -(q[j] + s[0]);
    const double zj = Minv[j] * rj;
    r[j] = rj; p[j] = zj; xt[j] = x[j];
    asum[0] += rj * zj;
    amax[0] = fmax(amax[0], fabs(rj));
  }
};
struct FEpiRhsB {    // r = -(q + [P|A'][x; v]), xt = x   (band preconditioner: z = K^-1 r follows)
  static constexpr int NSUM = 0, NMAX = 1;
  const double *q, *x;
  double *r, *xt;
  EPI_NO_PREFETCH
  __device__ __forceinline__ void row(int j, const double *s, const Pre &, double *, double *amax) const {
    const double rj = -(q[j] + s[0]);
    r[j] = rj; xt[j] = x[j];
    amax[0] = fmax(amax[0], fabs(rj));
  }
};
struct FEpiW {       // w = rho .* (A p)
  static constexpr int NSUM = 0, NMAX = 0;
  const double *rho;
  double *w;
  EPI_NO_PREFETCH
  __device__ __forceinline__ void row(int i, const double *s, const Pre &, double *, double *) const { w[i] = rho[i] * s[0]; }
};
struct FEpiKp {      // Kp = [P|A'][p; w] + sigma p ; p'Kp
  static constexpr int NSUM = 1, NMAX = 0;
  const double *p;
  double *Kp;
  double sigma;
  EPI_NO_PREFETCH
  __device__ __forceinline__ void row(int j, const double *s, const Pre &, double *asum, double *) const {
    const double pj = p[j];
    const double k = s[0] + sigma * pj;
    Kp[j] = k;
    asum[0] += pj * k;
  }
};
struct FEpiAdmm {    // z~ = A x~ ; z, y, Ax, v updates with box projection (EpiAdmm of engine.cu)
  static constexpr int NSUM = 0, NMAX = 0;
  const double *rho, *l, *u;
  double *z, *y, *Ax, *v, *dy;
  double alpha;
  int store;
  EPI_NO_PREFETCH
  __device__ __forceinline__ void row(int i, const double *s, const Pre &, double *, double *) const {
    const double zt = s[0], zo = z[i], yo = y[i], rh = rho[i];
    const double rinv = 1.0 / rh;
    const double zr = alpha * zt + (1.0 - alpha) * zo;
    const double zn = fmin(fmax(zr + rinv * yo, l[i]), u[i]);
    const double d = rh * (zr - zn);
    const double yn = yo + d;
    const double axn = alpha * zt + (1.0 - alpha) * Ax[i];
    z[i] = zn; y[i] = yn; Ax[i] = axn;
    v[i] = yn + rh * (axn - zn);
    if (store) dy[i] = d;
  }
};

// SpMV of the calling CTA over A (on_m = false) or M = [P | A'] with vectors in shared memory
template <int NV, bool SPLIT, int MODE, class Epi>
__device__ __forceinline__ void fused_spmv(const FusedArgs &a, bool on_m, const Gather &g0, const Gather &g1,
                                           Epi &e, double *asum, double *amax, int warp, int nwarps) {
  if (a.use_sell) spmv_sell_rows<NV, SPLIT, MODE, true>(on_m ? a.Ms : a.As, g0, g1, e, asum, amax, warp, nwarps);
  else spmv_direct_rows<NV, SPLIT, MODE, true>(on_m ? a.M : a.A, g0, g1, e, asum, amax, warp, nwarps);
}

__device__ __forceinline__ void admm_fused_body(const FusedArgs &a) {
  extern __shared__ __align__(16) double fsm[];
  __shared__ double scratch[32];
  const int n = a.n, m = a.m;
  double *x = fsm, *xt = x + n, *r = xt + n, *p = r + n, *Kp = p + n;      // 5 n
  double *z = Kp + n, *y = z + m, *v = y + m, *w = v + m, *Ax = w + m;     // 5 m
  double *t1 = Ax + m, *t2 = t1 + (a.use_band ? a.band.npad : 0);          // band scratch (2 npad)
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int warp = tid >> 5, nwarps = nthr >> 5;

  for (int j = tid; j < n; j += nthr) x[j] = a.x[j];
  for (int i = tid; i < m; i += nthr) { z[i] = a.z[i]; y[i] = a.y[i]; v[i] = a.v[i]; Ax[i] = a.Ax[i]; }
  __syncthreads();

  const double cap = a.use_cap ? a.S->tol_cap : DBL_MAX;
  long long tot_it = 0;
  int n_maxed = 0, max_it_seen = 0, n_break = 0, noncvx = 0, last_it = 0, last_conv = 0;
  const Gather none{nullptr, nullptr};

  for (int k = 0; k < a.count; ++k) {
    const int st = (a.store && k == a.count - 1) ? 1 : 0;
    // ---- rhs, r0, p0, x~ <- x
    double asum[1] = {0.0}, amax[1] = {-DBL_MAX};
    double rz, rn;
    if (a.use_band) {
      {
        FEpiRhsB e{a.q, x, r, xt};
        fused_spmv<1, false, 0>(a, true, Gather{x, v}, none, e, asum, amax, warp, nwarps);
      }
      rn = block_reduce<true>(amax[0], scratch);
      __syncthreads();
      band_apply(a.band, r, t1, t2);                 // z = K^-1 r0 (in t1, level-set order)
      double a_rz = 0.0;
      for (int j = tid; j < n; j += nthr) {
        const double zj = t1[__ldg(a.band.old2new + j)];
        p[j] = zj;
        a_rz += r[j] * zj;
      }
      rz = block_reduce<false>(a_rz, scratch);
    } else {
      {
        FEpiRhs e{a.q, a.Minv, x, r, p, xt};
        fused_spmv<1, false, 0>(a, true, Gather{x, v}, none, e, asum, amax, warp, nwarps);
      }
      rz = block_reduce<false>(asum[0], scratch);
      rn = block_reduce<true>(amax[0], scratch);
    }
    __syncthreads();
    double tol = a.rel * rn;
    if (a.use_cap) tol = fmin(tol, cap);
    tol = fmax(tol, a.absf);
    // ---- CG on K d = r0 (x~ accumulates x + d)
    int it = 0, conv = 0;
    for (;;) {
      conv = !(rn > tol);
      if (conv || it >= a.maxit) break;
      {
        FEpiW e{a.rho, w};
        fused_spmv<1, false, 0>(a, false, Gather{p, nullptr}, none, e, asum, amax, warp, nwarps);
      }
      __syncthreads();
      double pk[1] = {0.0};
      {
        FEpiKp e{p, Kp, a.sigma};
        fused_spmv<1, false, 0>(a, true, Gather{p, w}, none, e, pk, amax, warp, nwarps);
      }
      const double pKp = block_reduce<false>(pk[0], scratch);
      if (!(pKp > 0.0)) {                // negative curvature / breakdown: K is not positive definite
        if (rz > 0.0 && pKp <= 0.0) noncvx = 1;
        n_break += 1;
        break;
      }
      const double al = rz / pKp;
      double a_rz = 0.0, a_rn = -DBL_MAX;
      if (a.use_band) {
        for (int j = tid; j < n; j += nthr) {
          xt[j] += al * p[j];
          const double rj = r[j] - al * Kp[j];
          r[j] = rj;
          a_rn = fmax(a_rn, fabs(rj));
        }
        rn = block_reduce<true>(a_rn, scratch);
        ++it;
        __syncthreads();
        if (rn > tol && it < a.maxit) {              // another iteration follows: z = K^-1 r, new direction
          band_apply(a.band, r, t1, t2);
          for (int j = tid; j < n; j += nthr) a_rz += r[j] * t1[__ldg(a.band.old2new + j)];
          const double rz_new = block_reduce<false>(a_rz, scratch);
          const double beta = rz_new / rz;
          for (int j = tid; j < n; j += nthr) p[j] = t1[__ldg(a.band.old2new + j)] + beta * p[j];
          rz = rz_new;
          __syncthreads();
        }
        continue;
      }
      for (int j = tid; j < n; j += nthr) {
        xt[j] += al * p[j];
        const double rj = r[j] - al * Kp[j];
        r[j] = rj;
        a_rz += rj * (a.Minv[j] * rj);
        a_rn = fmax(a_rn, fabs(rj));
      }
      const double rz_new = block_reduce<false>(a_rz, scratch);
      rn = block_reduce<true>(a_rn, scratch);
      const double beta = rz_new / rz;
      for (int j = tid; j < n; j += nthr) p[j] = a.Minv[j] * r[j] + beta * p[j];
      rz = rz_new;
      ++it;
      __syncthreads();
    }
    tot_it += it;
    if (!conv) n_maxed += 1;
    if (it > max_it_seen) max_it_seen = it;
    last_it = it; last_conv = conv;
    // ---- x update, then z~ = A x~ fused with the z / y / Ax / v updates
    for (int j = tid; j < n; j += nthr) {
      const double xo = x[j];
      const double xn = a.alpha * xt[j] + (1.0 - a.alpha) * xo;
      x[j] = xn;
      if (st) a.dx[j] = xn - xo;
    }
    {
      FEpiAdmm e{a.rho, a.l, a.u, z, y, Ax, v, a.dy, a.alpha, st};
      fused_spmv<1, false, 0>(a, false, Gather{xt, nullptr}, none, e, asum, amax, warp, nwarps);
    }
    __syncthreads();
  }

  for (int j = tid; j < n; j += nthr) a.x[j] = x[j];
  for (int i = tid; i < m; i += nthr) { a.z[i] = z[i]; a.y[i] = y[i]; a.v[i] = v[i]; a.Ax[i] = Ax[i]; }
  if (tid == 0) {
    PcgScal *S = a.S;
    S->total_iters += tot_it;
    S->solves += a.count;
    S->maxed += n_maxed;
    if (max_it_seen > S->max_iters) S->max_iters = max_it_seen;
    S->breakdowns += n_break;
    if (noncvx) S->noncvx = 1;
    S->iters = last_it;
    S->conv = last_conv;
    S->done = 1;
  }
}

__global__ void __launch_bounds__(ADMM_FUSED_THREADS, 1) admm_fused_kernel(FusedArgs a) { admm_fused_body(a); }

// Coalesced launch (osqp_solve_batch): CTA b runs the iterations of problem b.
__global__ void __launch_bounds__(ADMM_FUSED_THREADS, 1) admm_fused_batch_kernel(const FusedArgs *args) {
  __shared__ FusedArgs a;
  if (threadIdx.x == 0) a = args[blockIdx.x];
  __syncthreads();
  admm_fused_body(a);
}
