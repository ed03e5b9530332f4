// pcg.cuh -- device-resident Jacobi-preconditioned conjugate gradient on the
// reduced KKT system  K x = rhs,  K = P + sigma I + A' diag(rho) A.
//
// Replaces the QDLDL factor/solve of reference
// lin_sys/direct/qdldl/qdldl_interface.c:53-103,341-376 (and qdldl.c:72-281).
// One CG iteration = 4 launches, no host synchronisation:
//   K_b  w  = rho .* (A p)                      stream SpMV over CSR(A)
//   K_c  Kp = [P | A'] [p ; w] + sigma p, p'Kp  stream SpMV over CSR(M)
//   K_d  alpha = rz / pKp; x += alpha p; r -= alpha Kp; rz', rr' partials
//   K_e  beta = rz'/rz; p = Minv r + beta p
// Scalars never leave the device: every CTA re-reduces the per-CTA partials of
// the previous kernel (fixed order -> identical value in every CTA, and
// bit-reproducible run to run).  Convergence sets a device flag that turns the
// remaining already-enqueued launches into no-ops.
#pragma once
#include "dev_mats.cuh"

struct PcgScal {          // lives in device memory
  double rz;              // r'z entering the current iteration
  double tol;             // residual target ||r||inf of the current solve
  double r0norm;          // ||r0||inf of the current solve
  double tol_cap;         // lambda * sqrt(pri_res * dua_res) of the last residual check (ADMM solves)
  int done;               // 1: remaining launches of this solve are no-ops
  int conv;               // 1: residual target met
  int iters;              // CG iterations of the current solve
  int noncvx;             // sticky: p'Kp <= 0 or non-positive Jacobi diagonal seen
  long long total_iters;  // since creation
  int solves;
  int maxed;              // solves that ran out of enqueued iterations
  int max_iters;          // largest `iters` since the host last reset it
  int breakdowns;         // solves stopped by p'Kp <= 0 / NaN
  double rzb[2];          // r'z of iterations it (slot it & 1) and it - 1: beta inside K_b (EpiW::fuse)
  double alpha_last;      // step length of the last APPLIED CG iteration of the current solve
  int applied;            // CG iterations of the current solve whose update was applied
};

struct PcgDev {
  DevMats *W = nullptr;
  int n = 0, m = 0;
  double sigma = 0.0;
  double *rho = nullptr;    // m   (owned by the engine or the plugin instance)
  double *Minv = nullptr;   // n   inverse Jacobi diagonal
  double *r = nullptr, *p = nullptr, *Kp = nullptr;   // n
  double *z = nullptr;      // n: M^-1 r, stored by K_d for the ADMM solves (K_e folded into K_b: EpiW::fuse)
  double *w = nullptr;      // m
  double *partRZ[2] = {nullptr, nullptr};   // [rz, rr] pairs, 2 * MAX blocks each
  double *partPKP = nullptr;
  PcgScal *S = nullptr;
  double rel_tol = 1e-2, abs_floor = 1e-30;
  int use_cap = 0;          // ADMM solves: also cap the target by PcgScal::tol_cap
  int max_iter = 500;
  int grid_n = 1;           // CTAs of the elementwise n-kernels
  int gridA = 1, gridM = 1; // CTAs of the stream kernels
  cudaStream_t st = 0;
};

// ------------------------------------------------------------- epilogues
struct EpiPlain {                      // no fused work
  static constexpr int NSUM = 0, NMAX = 0;
  __device__ bool begin(double *) { return true; }
  EPI_NO_PREFETCH
  __device__ void row(int, const double *, const Pre &, double *, double *) const {}
};

// K_b: w = rho .* (A p), preceded by the convergence test of iteration `it`
struct EpiW {
  static constexpr int NSUM = 0, NMAX = 0;
  PcgScal *S;
  const double *part;     // [r'z, ||r||inf] pairs produced before this iteration
  int nparts, it, maxit;
  double rel, absf;       // target = max(absf, min(rel * ||r0||inf, tol_cap if use_cap))
  int use_cap;
  const double *rho;
  double *w;
  // ADMM solves only (else null): zacc accumulates A (x~ - x) = sum_k alpha_k A p_k, with
  // A p_k = w_k ./ rho read back just before w is overwritten, so that z~ = A x~ needs no SpMV
  // of its own (reference src/auxil.c:177-183 gets z~ from the KKT solve for the same reason)
  double *zacc;
  // ADMM solves: K_e (p = z + beta p) is folded in here.  This kernel gathers z = M^-1 r (stored by K_d)
  // instead of p and uses  A p = A z + beta A p_old,  i.e.  w = rho .* (A z) + beta w_old;  the
  // direction update itself, which only the LATER kernels of the iteration read, is done by begin(),
  // every CTA taking a share of p.  One launch less per CG iteration (7.6 us of ~125 at n = 1e5).
  int fuse, n;
  const double *Minv, *r;
  double *p;
  double alpha_prev, beta;      // filled by begin()
  __device__ bool begin(double *scratch) {
    if (S->done) return false;
    const double rz = reduce_partials_sum(part, nparts, 2, scratch);
    const double rn = reduce_partials_max(part + 1, nparts, 2, scratch);
    // r'z of the previous iteration sits in the other slot: no race with CTA 0 writing this one
    beta = (fuse && it > 0) ? rz / S->rzb[(it - 1) & 1] : 0.0;
    double tol = S->tol;
    if (it == 0) {
      tol = rel * rn;
      if (use_cap) tol = fmin(tol, S->tol_cap);
      tol = fmax(tol, absf);
    }
    const bool conv = !(rn > tol);
    const bool stop = conv || it >= maxit;
    alpha_prev = S->alpha_last;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      if (it == 0) { S->tol = tol; S->r0norm = rn; }
      if (stop) { S->done = 1; S->conv = conv ? 1 : 0; }
      else { S->rz = rz; S->rzb[it & 1] = rz; S->iters = it + 1; }
    }
    if (stop) return false;
    if (fuse && it > 0) {
      const int stride = gridDim.x * blockDim.x;
      for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride) p[j] = Minv[j] * r[j] + beta * p[j];
    }
    return true;
  }
  struct Pre { double rho, w_old; };
  __device__ __forceinline__ Pre load(int i) const {
    return Pre{rho[i], (zacc && it > 0) ? w[i] : 0.0};
  }
  __device__ void row(int i, const double *s, const Pre &pre, double *, double *) const {
    if (zacc && it > 0) zacc[i] += alpha_prev * (pre.w_old / pre.rho);
    w[i] = fuse ? fma(beta, pre.w_old, pre.rho * s[0]) : pre.rho * s[0];
  }
};

// w = rho .* (A x0) outside the CG loop (initial residual of a general solve)
struct EpiWPlain {
  static constexpr int NSUM = 0, NMAX = 0;
  const double *rho;
  double *w;
  __device__ bool begin(double *) { return true; }
  struct Pre { double rho; };
  __device__ __forceinline__ Pre load(int i) const { return Pre{rho[i]}; }
  __device__ void row(int i, const double *s, const Pre &pre, double *, double *) const { w[i] = pre.rho * s[0]; }
};

// K_c: Kp = M [p ; w] + sigma p ; partial of p'Kp
struct EpiKp {
  static constexpr int NSUM = 1, NMAX = 0;
  const PcgScal *S;
  const double *p;
  double *Kp;
  double sigma;
  __device__ bool begin(double *) { return !S->done; }
  struct Pre { double p; };
  __device__ __forceinline__ Pre load(int j) const { return Pre{p[j]}; }
  __device__ void row(int j, const double *s, const Pre &pre, double *asum, double *) const {
    const double pj = pre.p;
    const double k = s[0] + sigma * pj;
    Kp[j] = k;
    asum[0] += pj * k;
  }
};

template <> struct epi_tile_ok<EpiW> { static constexpr bool value = true; };
template <> struct epi_tile_ok<EpiWPlain> { static constexpr bool value = true; };
template <> struct epi_tile_ok<EpiKp> { static constexpr bool value = true; };

// general start: r = b - (M [x0 ; w] + sigma x0), p = Minv r, partials [r'z, ||r||inf]
struct EpiInitGeneral {
  static constexpr int NSUM = 1, NMAX = 1;
  PcgScal *S;
  const double *b, *x0, *Minv;
  double *r, *p;
  double sigma;
  int x0_is_zero;
  __device__ bool begin(double *) {
    if (blockIdx.x == 0 && threadIdx.x == 0) { S->done = 0; S->conv = 0; S->iters = 0; S->applied = 0; }
    return true;
  }
  struct Pre { double b, x0, Minv; };
  __device__ __forceinline__ Pre load(int j) const { return Pre{b[j], x0_is_zero ? 0.0 : x0[j], Minv[j]}; }
  __device__ void row(int j, const double *s, const Pre &pre, double *asum, double *amax) const {
    const double rj = x0_is_zero ? pre.b : pre.b - (s[0] + sigma * pre.x0);
    const double zj = pre.Minv * rj;
    r[j] = rj;
    p[j] = zj;
    asum[0] += rj * zj;
    amax[0] = fmax(amax[0], fabs(rj));
  }
};

template <> struct epi_tile_ok<EpiInitGeneral> { static constexpr bool value = true; };

// Jacobi: Minv = 1 / (P_jj + sigma + sum_i rho_i A_ij^2)   (stream MODE 1)
struct EpiJacobi {
  static constexpr int NSUM = 0, NMAX = 0;
  PcgScal *S;
  double *Minv;
  double sigma;
  __device__ bool begin(double *) { return true; }
  EPI_NO_PREFETCH
  __device__ void row(int j, const double *s, const Pre &, double *, double *) const {
    const double d = s[0] + sigma;
    if (!(d > 0.0)) S->noncvx = 1;
    Minv[j] = 1.0 / d;
  }
};

// ------------------------------------------------- elementwise CG kernels
// K_d
// Both kernels are latency-bound at n ~ 1e5 (one element per thread): the element's inputs are
// fetched BEFORE the scalar reduction so that the two round trips overlap.
__global__ void __launch_bounds__(OSQP_CUDA_THREADS)
pcg_update_kernel(PcgScal *S, const double *partPKP, int nparts_pkp, int n, double *x, double *r,
                  const double *p, const double *Kp, const double *Minv, double *part_out, double *zout) {
  __shared__ double scratch[32];
  const int j0 = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
  double x0 = 0.0, p0 = 0.0, r0 = 0.0, k0 = 0.0, m0 = 0.0;
  if (j0 < n) { x0 = x[j0]; p0 = p[j0]; r0 = r[j0]; k0 = Kp[j0]; m0 = Minv[j0]; }
  if (S->done) return;
  const double pKp = reduce_partials_sum(partPKP, nparts_pkp, 1, scratch);
  const double rz = S->rz;
  if (!(pKp > 0.0)) {       // negative curvature (or breakdown): K is not positive definite
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      if (rz > 0.0 && pKp <= 0.0) S->noncvx = 1;
      S->breakdowns += 1;
      S->done = 1;
    }
    return;
  }
  const double alpha = rz / pKp;
  if (blockIdx.x == 0 && threadIdx.x == 0) { S->alpha_last = alpha; S->applied += 1; }
  double a_rz = 0.0, a_rn = 0.0;
  if (j0 < n) {
    x[j0] = x0 + alpha * p0;
    const double rj = m0 == 0.0 ? 0.0 : r0 - alpha * k0;      // Minv == 0: eliminated slack (slack.cuh)
    r[j0] = rj;
    if (zout) zout[j0] = m0 * rj;
    a_rz += rj * (m0 * rj);
    a_rn = fmax(a_rn, fabs(rj));
  }
  for (int j = j0 + stride; j < n; j += stride) {
    x[j] += alpha * p[j];
    const double mj = Minv[j];
    const double rj = mj == 0.0 ? 0.0 : r[j] - alpha * Kp[j];
    r[j] = rj;
    if (zout) zout[j] = mj * rj;
    a_rz += rj * (mj * rj);
    a_rn = fmax(a_rn, fabs(rj));
  }
  const double t0 = block_reduce<false>(a_rz, scratch);
  const double t1 = block_reduce<true>(a_rn, scratch);
  if (threadIdx.x == 0) {
    part_out[2 * blockIdx.x] = t0;
    part_out[2 * blockIdx.x + 1] = t1;
  }
}
// K_e
__global__ void __launch_bounds__(OSQP_CUDA_THREADS)
pcg_dir_kernel(const PcgScal *S, const double *part_new, int nparts, int n, double *p,
               const double *r, const double *Minv) {
  __shared__ double scratch[32];
  const int j0 = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
  double p0 = 0.0, r0 = 0.0, m0 = 0.0;
  if (j0 < n) { p0 = p[j0]; r0 = r[j0]; m0 = Minv[j0]; }
  if (S->done) return;
  const double rz_new = reduce_partials_sum(part_new, nparts, 2, scratch);
  const double beta = rz_new / S->rz;
  if (j0 < n) p[j0] = m0 * r0 + beta * p0;
  for (int j = j0 + stride; j < n; j += stride) p[j] = Minv[j] * r[j] + beta * p[j];
}
// closes a solve: statistics + "ran out of enqueued iterations" detection
__device__ __forceinline__ void pcg_close(PcgScal *S, const double *part_last, int nparts,
                                          double *scratch) {
  // called by ONE CTA (all its threads)
  double rn = 0.0;
  const int done = S->done;
  if (!done) rn = reduce_partials_max(part_last + 1, nparts, 2, scratch);
  if (threadIdx.x == 0) {
    int conv = S->conv;
    if (!done) conv = !(rn > S->tol);
    S->total_iters += S->iters;
    S->solves += 1;
    if (!conv) S->maxed += 1;
    if (S->iters > S->max_iters) S->max_iters = S->iters;
    S->done = 1;
    S->conv = conv;
  }
}
__global__ void __launch_bounds__(OSQP_CUDA_THREADS)
pcg_close_kernel(PcgScal *S, const double *part_last, int nparts) {
  __shared__ double scratch[32];
  pcg_close(S, part_last, nparts, scratch);
}
