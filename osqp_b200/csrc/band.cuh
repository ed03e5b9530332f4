// band.cuh -- exact block-tridiagonal preconditioner for SMALL problems whose reduced KKT matrix
//   K = P + sigma I + A' diag(rho) A
// is banded under a level-set ordering (MPC problems of BASELINE config 4: horizon x (states +
// inputs), where Jacobi-preconditioned CG needs ~120 iterations per solve because the equality rows
// carry rho x 1e3 -- reference src/auxil.c:90, include/constants.h:59,75).  SURVEY 8(f) rank 3.
//
// Replaces, for those problems, the role of the reference's LDL' factorisation of the KKT matrix
// (lin_sys/direct/qdldl/qdldl_interface.c:53-103, refactored on every rho update :396-410) with a
// factorisation that maps to one thread block:
//   * setup (host, once per sparsity pattern -- cached, a batch of 4096 MPC problems shares one plan):
//     the graph of K = pattern(P) + pattern(A'A); breadth-first LEVEL SETS from a pseudo-peripheral
//     vertex.  Edges only join adjacent levels, so with the variables ordered level by level K is
//     block tridiagonal BY CONSTRUCTION, with blocks of B = (largest level, rounded up to 8) <= 64.
//     (MPC: 51 levels of <= 49 variables.  A random QP has levels of ~n and does not qualify.)
//   * every rho / sigma / matrix update (device, k_band_build + k_band_factor, one CTA):
//     the diagonal blocks D_k and sub-diagonal blocks E_k of K from the CSR arrays, then the block
//     Cholesky  L_kk L_kk' = D_k - L_k,k-1 L_k,k-1',  L_k,k-1 = E_k L_k-1,k-1^-T,  kept as
//       G_k = L_kk^-1,   H_k = G_k L_k,k-1,   JT_k-1 = (L_k,k-1 G_k-1)'
//     so that a solve is nothing but small dense matrix-vector products:
//       forward   y_k = G_k r_k - H_k y_k-1          backward   x_k = G_k' y_k - JT_k x_k+1
//     G_k r_k and G_k' y_k for all k in parallel, the two recurrences one block per step.
//     A non-positive pivot means K is not positive definite: the same condition the reference
//     reads off the LDL' inertia (qdldl_interface.c:93-98) -> PcgScal::noncvx.
//   * inside the fused single-CTA ADMM kernel (admm_fused.cuh) z = K^-1 r replaces z = Minv r: the
//     CG converges in one iteration (the second only confirms the residual), and the ADMM
//     iteration count is that of the reference's exact solves.
// Deterministic: fixed summation orders everywhere.
#pragma once
#include "dev_mats.cuh"
#include <algorithm>
#include <mutex>
#include <unordered_map>

#define BAND_BMAX 64
#define BAND_THREADS 1024
#define BAND_MAX_N 4096
#define BAND_MAX_ROW 256          // longest row of A the pattern pass accepts (pairs grow with its square)

struct BandView {                 // device pointers, all in the level-set ordering
  int B, nl, npad;
  const int *new2old;             // [npad], -1 = padding
  const int *old2new;             // [n]
  const double *G, *H, *JT;       // [nl][B][B] row-major; H_0 and JT_(nl-1) unused
};
struct BandDev {
  bool ok = false;
  int n = 0, B = 0, nl = 0, npad = 0;
  int *new2old = nullptr, *old2new = nullptr;
  // G and H hold the blocks D_k, E_k of K between k_band_build and k_band_factor, which replaces
  // them level by level (D_k and E_k are read before G_k and H_k are written)
  double *G = nullptr, *H = nullptr, *JT = nullptr;
};
static inline BandView band_view_of(const BandDev &b) {
  return BandView{b.B, b.nl, b.npad, b.new2old, b.old2new, b.G, b.H, b.JT};
}
static inline size_t band_apply_smem(const BandDev &b) { return b.ok ? sizeof(double) * 2 * (size_t)b.npad : 0; }

// ------------------------------------------------------------------ host: level-set plan
struct BandPlan { bool ok = false; int n = 0, B = 0, nl = 0; std::vector<int> new2old, old2new; };

static unsigned long long band_hash(const csc *P, const csc *A) {
  unsigned long long h = 1469598103934665603ull;
  auto mix = [&](const c_int *p, long long cnt) {
    for (long long i = 0; i < cnt; ++i) { h ^= (unsigned long long)p[i]; h *= 1099511628211ull; }
  };
  const c_int dims[4] = {P->n, A->m, P->p[P->n], A->p ? A->p[A->n] : 0};
  mix(dims, 4);
  mix(P->p, P->n + 1); mix(P->i, P->p[P->n]);
  if (A->p) { mix(A->p, A->n + 1); mix(A->i, A->p[A->n]); }
  return h;
}

static void band_plan_compute(const csc *P, const csc *A, BandPlan &out) {
  const int n = (int)P->n, m = (int)A->m;
  out = BandPlan();
  out.n = n;
  if (n < 64 || n > BAND_MAX_N) return;
  const long long nnzA = A->p ? A->p[A->n] : 0;
  // rows of A (CSR pattern by counting sort)
  std::vector<int> rp(m + 1, 0), rc((size_t)nnzA);
  for (long long e = 0; e < nnzA; ++e) rp[A->i[e] + 1]++;
  for (int r = 0; r < m; ++r) {
    if (rp[r + 1] > BAND_MAX_ROW) return;
    rp[r + 1] += rp[r];
  }
  {
    std::vector<int> fill(rp.begin(), rp.end() - 1);
    for (int j = 0; j < n; ++j)
      for (c_int e = A->p[j]; e < A->p[j + 1]; ++e) rc[fill[A->i[e]]++] = j;
  }
  // adjacency of K as a bit matrix
  const int words = (n + 63) / 64;
  std::vector<unsigned long long> bits((size_t)n * words, 0ull);
  auto set = [&](int a, int b) { bits[(size_t)a * words + (b >> 6)] |= 1ull << (b & 63); };
  for (int j = 0; j < n; ++j)
    for (c_int e = P->p[j]; e < P->p[j + 1]; ++e) { set((int)P->i[e], j); set(j, (int)P->i[e]); }
  for (int r = 0; r < m; ++r)
    for (int a = rp[r]; a < rp[r + 1]; ++a)
      for (int b = rp[r]; b < rp[r + 1]; ++b) set(rc[a], rc[b]);
  std::vector<int> ap(n + 1, 0), adj;
  for (int v = 0; v < n; ++v) {
    for (int w = 0; w < words; ++w) {
      unsigned long long t = bits[(size_t)v * words + w];
      while (t) { const int b = __builtin_ctzll(t); t &= t - 1; const int u = w * 64 + b; if (u != v) adj.push_back(u); }
    }
    ap[v + 1] = (int)adj.size();
  }
  // breadth-first levels; a new component starts in the level after the last one (no edges between)
  std::vector<int> lvl(n), order;
  std::vector<int> lstart;
  auto bfs = [&](int start) {
    std::fill(lvl.begin(), lvl.end(), -1);
    order.clear(); lstart.clear();
    std::vector<int> cur(1, start), nxt;
    lvl[start] = 0;
    int k = 0, seen = 0, scan = 0;
    for (;;) {
      lstart.push_back((int)order.size());
      std::sort(cur.begin(), cur.end());
      for (int v : cur) order.push_back(v);
      seen += (int)cur.size();
      nxt.clear();
      for (int v : cur)
        for (int e = ap[v]; e < ap[v + 1]; ++e) { const int u = adj[e]; if (lvl[u] < 0) { lvl[u] = k + 1; nxt.push_back(u); } }
      if (nxt.empty()) {
        if (seen == n) break;
        while (lvl[scan] >= 0) ++scan;
        lvl[scan] = k + 1; nxt.push_back(scan);
      }
      cur.swap(nxt); ++k;
    }
    lstart.push_back((int)order.size());
  };
  int start = 0;
  for (int v = 1; v < n; ++v) if (ap[v + 1] - ap[v] < ap[start + 1] - ap[start]) start = v;
  bfs(start);
  for (int sweep = 0; sweep < 3; ++sweep) {          // pseudo-peripheral vertex (George-Liu)
    const int depth = (int)lstart.size() - 1;
    int cand = order[lstart[depth - 1]];
    for (int t = lstart[depth - 1]; t < lstart[depth]; ++t)
      if (ap[order[t] + 1] - ap[order[t]] < ap[cand + 1] - ap[cand]) cand = order[t];
    std::vector<int> keep_order = order, keep_lstart = lstart;
    bfs(cand);
    if ((int)lstart.size() - 1 <= depth) { order.swap(keep_order); lstart.swap(keep_lstart); break; }
  }
  const int nl = (int)lstart.size() - 1;
  int wmax = 0;
  for (int k = 0; k < nl; ++k) wmax = std::max(wmax, lstart[k + 1] - lstart[k]);
  if (wmax > BAND_BMAX || nl < 4) return;
  const int B = std::max(8, (wmax + 7) / 8 * 8);
  out.B = B; out.nl = nl;
  out.new2old.assign((size_t)nl * B, -1);
  out.old2new.assign(n, -1);
  for (int k = 0; k < nl; ++k)
    for (int t = lstart[k]; t < lstart[k + 1]; ++t) {
      const int pos = k * B + (t - lstart[k]);
      out.new2old[pos] = order[t]; out.old2new[order[t]] = pos;
    }
  out.ok = true;
}

// cached by sparsity pattern: the 4096 problems of a batch differ in their bounds only
static bool band_plan(const csc *P, const csc *A, BandPlan &out) {
  static std::mutex mu;
  static std::unordered_map<unsigned long long, BandPlan> cache;
  const unsigned long long h = band_hash(P, A);
  {
    std::lock_guard<std::mutex> lock(mu);
    auto it = cache.find(h);
    if (it != cache.end() && it->second.n == (int)P->n) { out = it->second; return out.ok; }
  }
  band_plan_compute(P, A, out);
  std::lock_guard<std::mutex> lock(mu);
  if (cache.size() > 64) cache.clear();
  cache[h] = out;
  return out.ok;
}

// ------------------------------------------------------------------ device: blocks of K
// D <- 0 with ones on the diagonal of the padding slots, E <- 0
__global__ void k_band_init(const int *new2old, int B, int npad, double *D, double *E) {
  const size_t total = (size_t)npad * B;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int pos = (int)(i / B), t = (int)(i % B);
    D[i] = (new2old[pos] < 0 && t == pos % B) ? 1.0 : 0.0;
    E[i] = 0.0;
  }
}
// One warp per variable i (old numbering): row i of K restricted to its own level k and the
// two neighbours (all of it, by construction), accumulated in a fixed order:
//   K_ij = P_ij + sigma [i == j] + sum_r A_ri rho_r A_rj
// from row i of M = [Pf | A'] (P entries, then the A' entries (r, A_ri)) and row r of A.
__global__ void __launch_bounds__(256)
k_band_build(CsrView M, CsrView A, const double *__restrict__ rho, double sigma, int n, int B,
             const int *__restrict__ old2new, double *D, double *E) {
  extern __shared__ double band_buf[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  double *buf = band_buf + (size_t)warp * 3 * B;
  for (int i = blockIdx.x * nw + warp; i < n; i += gridDim.x * nw) {
    const int pos = __ldg(old2new + i), k = pos / B, s = pos - k * B;
    for (int t = lane; t < 3 * B; t += 32) buf[t] = 0.0;
    __syncwarp();
    const int e0 = __ldg(M.rowptr + i), e1 = __ldg(M.rowptr + i + 1);
    for (int e = e0; e < e1; ++e) {
      const int c = __ldg(M.col + e);
      const double v = __ldg(M.val + e);
      if (c < n) {
        if (lane == 0) { const int pj = __ldg(old2new + c); buf[(pj / B - k + 1) * B + pj % B] += v; }
      } else {
        const int r = c - n;
        const double f = __ldg(rho + r) * v;
        for (int t = __ldg(A.rowptr + r) + lane; t < __ldg(A.rowptr + r + 1); t += 32) {
          const int pj = __ldg(old2new + __ldg(A.col + t));
          buf[(pj / B - k + 1) * B + pj % B] += f * __ldg(A.val + t);      // distinct columns: no conflicts
        }
      }
      __syncwarp();
    }
    if (lane == 0) buf[B + s] += sigma;
    __syncwarp();
    for (int t = lane; t < B; t += 32) {
      D[(size_t)pos * B + t] = buf[B + t];
      if (k > 0) E[(size_t)pos * B + t] = buf[t];
    }
    __syncwarp();
  }
}

// ------------------------------------------------------------------ device: block Cholesky
// One CTA.  Shared memory: five (B x (B + 1)) matrices.
__device__ __forceinline__ void band_gemm_nt(const double *X, const double *Y, double *Z, int B, int ld, double sign,
                                             bool accumulate) {
  // Z[a][b] (+)= sign * sum_t X[a][t] Y[b][t]
  for (int o = threadIdx.x; o < B * B; o += blockDim.x) {
    const int a = o / B, b = o - a * B;
    double acc = 0.0;
    for (int t = 0; t < B; ++t) acc = fma(X[a * ld + t], Y[b * ld + t], acc);
    Z[a * ld + b] = (accumulate ? Z[a * ld + b] : 0.0) + sign * acc;
  }
}
__device__ __forceinline__ void band_gemm_nn(const double *X, const double *Y, double *Z, int B, int ld) {
  // Z[a][b] = sum_t X[a][t] Y[t][b]
  for (int o = threadIdx.x; o < B * B; o += blockDim.x) {
    const int a = o / B, b = o - a * B;
    double acc = 0.0;
    for (int t = 0; t < B; ++t) acc = fma(X[a * ld + t], Y[t * ld + b], acc);
    Z[a * ld + b] = acc;
  }
}
__global__ void __launch_bounds__(BAND_THREADS, 1)
k_band_factor(int B, int nl, double *G, double *H, double *JT, int *noncvx) {
  const double *D = G, *E = H;                               // in: blocks of K; out: the factor, in place
  extern __shared__ double band_sm[];
  const int ld = B + 1, tid = threadIdx.x, nthr = blockDim.x;
  double *Sd = band_sm, *Gk = Sd + B * ld, *Gp = Gk + B * ld, *Lk = Gp + B * ld, *W = Lk + B * ld;
  for (int k = 0; k < nl; ++k) {
    const double *Dk = D + (size_t)k * B * B, *Ek = E + (size_t)k * B * B;
    for (int o = tid; o < B * B; o += nthr) { const int a = o / B, b = o - a * B; Sd[a * ld + b] = Dk[o]; if (k) W[a * ld + b] = Ek[o]; }
    __syncthreads();
    if (k) {
      band_gemm_nt(W, Gp, Lk, B, ld, 1.0, false);           // L_k,k-1 = E_k G_k-1'
      __syncthreads();
      band_gemm_nt(Lk, Lk, Sd, B, ld, -1.0, true);          // S = D_k - L_k,k-1 L_k,k-1'
      __syncthreads();
    }
    // ---- Cholesky of S in place (lower triangle), right-looking
    for (int j = 0; j < B; ++j) {
      double d = Sd[j * ld + j];
      if (!(d > 0.0)) { if (tid == 0) *noncvx = 1; d = 1.0; }   // K is not positive definite
      const double ljj = sqrt(d);
      __syncthreads();
      for (int i = j + tid; i < B; i += nthr) Sd[i * ld + j] = (i == j) ? ljj : Sd[i * ld + j] / ljj;
      __syncthreads();
      const int rem = B - j - 1;
      for (int o = tid; o < rem * rem; o += nthr) {
        const int i = j + 1 + o / rem, c = j + 1 + o % rem;
        if (c <= i) Sd[i * ld + c] -= Sd[i * ld + j] * Sd[c * ld + j];
      }
      __syncthreads();
    }
    // ---- G_k = L_kk^-1 (lower triangular), row by row: 16 threads share the dot product of one column
    for (int o = tid; o < B * B; o += nthr) { const int a = o / B, b = o - a * B; Gk[a * ld + b] = 0.0; }
    __syncthreads();
    {
      const int c = tid >> 4, ts = tid & 15;
      for (int i = 0; i < B; ++i) {
        double acc = 0.0;
        if (c < B && c <= i)
          for (int t = c + ts; t < i; t += 16) acc = fma(Sd[i * ld + t], Gk[t * ld + c], acc);
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o, 16);
        if (c < B && c <= i && ts == 0) Gk[i * ld + c] = ((i == c ? 1.0 : 0.0) - acc) / Sd[i * ld + i];
        __syncthreads();
      }
    }
    for (int o = tid; o < B * B; o += nthr) { const int a = o / B, b = o - a * B; G[(size_t)k * B * B + o] = Gk[a * ld + b]; }
    if (k) {
      band_gemm_nn(Gk, Lk, W, B, ld);                       // H_k = G_k L_k,k-1
      __syncthreads();
      for (int o = tid; o < B * B; o += nthr) { const int a = o / B, b = o - a * B; H[(size_t)k * B * B + o] = W[a * ld + b]; }
      __syncthreads();
      band_gemm_nn(Lk, Gp, W, B, ld);                       // J_k-1 = L_k,k-1 G_k-1, stored transposed
      __syncthreads();
      for (int o = tid; o < B * B; o += nthr) { const int a = o / B, b = o - a * B; JT[(size_t)(k - 1) * B * B + o] = W[b * ld + a]; }
    }
    __syncthreads();
    for (int o = tid; o < B * ld; o += nthr) Gp[o] = Gk[o];
    __syncthreads();
  }
}

// ------------------------------------------------------------------ device: z = K^-1 r (one CTA)
// r: n doubles (shared memory, old numbering).  On return t1[new position] holds z: z_j is
// t1[old2new[j]].  t1, t2: npad doubles of shared memory each.  All threads of the CTA call.
__device__ __forceinline__ void band_apply(const BandView &b, const double *r, double *t1, double *t2) {
  const int B = b.B, nl = b.nl, tid = threadIdx.x, nthr = blockDim.x;
  const int lane = tid & 31, warp = tid >> 5, nw = nthr >> 5;
  const size_t BB = (size_t)B * B;
  for (int pos = tid; pos < b.npad; pos += nthr) { const int o = __ldg(b.new2old + pos); t1[pos] = o >= 0 ? r[o] : 0.0; }
  __syncthreads();
  // c = G_k r_k, every row at once: a warp per row, lanes over the (lower triangular) columns
  for (int row = warp; row < b.npad; row += nw) {
    const int k = row / B, s = row - k * B;
    const double *g = b.G + (size_t)k * BB + (size_t)s * B;
    double acc = 0.0;
    for (int t = lane; t <= s; t += 32) acc = fma(__ldg(g + t), t1[k * B + t], acc);
    acc = warp_sum(acc);
    if (lane == 0) t2[row] = acc;
  }
  __syncthreads();
  // forward recurrence y_k = c_k - H_k y_k-1, in place in t2; the rows of H_k+1 are fetched a step ahead
  {
    double h0 = 0.0, h1 = 0.0, g0 = 0.0, g1 = 0.0;
    auto fetch = [&](const double *Mx, int k, int row, double &a0, double &a1) {
      a0 = a1 = 0.0;
      if (k >= 0 && k < nl && row < B) {
        const double *p = Mx + (size_t)k * BB + (size_t)row * B;
        if (lane < B) a0 = __ldg(p + lane);
        if (lane + 32 < B) a1 = __ldg(p + lane + 32);
      }
    };
    // a warp owns rows warp and warp + 32 of every block
    fetch(b.H, 1, warp, h0, h1); fetch(b.H, 1, warp + 32, g0, g1);
    for (int k = 1; k < nl; ++k) {
      const double a0 = h0, a1 = h1, c0 = g0, c1 = g1;
      fetch(b.H, k + 1, warp, h0, h1); fetch(b.H, k + 1, warp + 32, g0, g1);
      const double *y = t2 + (k - 1) * B;
      double s0 = 0.0, s1 = 0.0;
      if (lane < B) { s0 = a0 * y[lane]; s1 = c0 * y[lane]; }
      if (lane + 32 < B) { s0 = fma(a1, y[lane + 32], s0); s1 = fma(c1, y[lane + 32], s1); }
      s0 = warp_sum(s0); s1 = warp_sum(s1);
      if (lane == 0) {
        if (warp < B) t2[k * B + warp] -= s0;
        if (warp + 32 < B) t2[k * B + warp + 32] -= s1;
      }
      __syncthreads();
    }
  }
  // d = G_k' y_k, every block at once: a warp per block, lanes over the columns, rows in order
  for (int k = warp; k < nl; k += nw) {
    const double *g = b.G + (size_t)k * BB;
    const double *y = t2 + k * B;
    double d0 = 0.0, d1 = 0.0;
    for (int t = 0; t < B; ++t) {
      const double yt = y[t];
      if (lane < B && lane <= t) d0 = fma(__ldg(g + (size_t)t * B + lane), yt, d0);
      if (lane + 32 < B && lane + 32 <= t) d1 = fma(__ldg(g + (size_t)t * B + lane + 32), yt, d1);
    }
    if (lane < B) t1[k * B + lane] = d0;
    if (lane + 32 < B) t1[k * B + lane + 32] = d1;
  }
  __syncthreads();
  // backward recurrence x_k = d_k - JT_k x_k+1, in place in t1
  {
    double h0 = 0.0, h1 = 0.0, g0 = 0.0, g1 = 0.0;
    auto fetch = [&](int k, int row, double &a0, double &a1) {
      a0 = a1 = 0.0;
      if (k >= 0 && k < nl - 1 && row < B) {
        const double *p = b.JT + (size_t)k * BB + (size_t)row * B;
        if (lane < B) a0 = __ldg(p + lane);
        if (lane + 32 < B) a1 = __ldg(p + lane + 32);
      }
    };
    fetch(nl - 2, warp, h0, h1); fetch(nl - 2, warp + 32, g0, g1);
    for (int k = nl - 2; k >= 0; --k) {
      const double a0 = h0, a1 = h1, c0 = g0, c1 = g1;
      fetch(k - 1, warp, h0, h1); fetch(k - 1, warp + 32, g0, g1);
      const double *xn = t1 + (k + 1) * B;
      double s0 = 0.0, s1 = 0.0;
      if (lane < B) { s0 = a0 * xn[lane]; s1 = c0 * xn[lane]; }
      if (lane + 32 < B) { s0 = fma(a1, xn[lane + 32], s0); s1 = fma(c1, xn[lane + 32], s1); }
      s0 = warp_sum(s0); s1 = warp_sum(s1);
      if (lane == 0) {
        if (warp < B) t1[k * B + warp] -= s0;
        if (warp + 32 < B) t1[k * B + warp + 32] -= s1;
      }
      __syncthreads();
    }
  }
  __syncthreads();
}

// ------------------------------------------------------------------ host: allocation, refactorisation
static void band_free(BandDev &b) {
  dev_free(b.new2old); dev_free(b.old2new); dev_free(b.G); dev_free(b.H); dev_free(b.JT);
  b.ok = false;
}
// plan (cached per pattern) + device buffers; b.ok = false when the problem does not qualify
static int band_setup(BandDev &b, const csc *P, const csc *A, cudaStream_t st) {
  band_free(b);
  const char *env = getenv("OSQP_B200_BAND");
  if (env && env[0] == '0') return 0;
  BandPlan plan;
  if (!band_plan(P, A, plan)) return 0;
  b.n = plan.n; b.B = plan.B; b.nl = plan.nl; b.npad = plan.nl * plan.B;
  const size_t blocks = (size_t)b.nl * b.B * b.B;
  if (dev_alloc(&b.new2old, (size_t)b.npad) || dev_alloc(&b.old2new, (size_t)b.n) || dev_alloc(&b.G, blocks) ||
      dev_alloc(&b.H, blocks) || dev_alloc(&b.JT, blocks)) {
    band_free(b);
    return 1;
  }
  if (cudaMemcpyAsync(b.new2old, plan.new2old.data(), sizeof(int) * b.npad, cudaMemcpyHostToDevice, st) != cudaSuccess ||
      cudaMemcpyAsync(b.old2new, plan.old2new.data(), sizeof(int) * b.n, cudaMemcpyHostToDevice, st) != cudaSuccess ||
      cudaStreamSynchronize(st) != cudaSuccess) { band_free(b); return 1; }
  cudaMemsetAsync(b.JT, 0, sizeof(double) * blocks, st);
  b.ok = true;
  return 0;
}
// K's blocks from the current device matrices, rho and sigma, then the block Cholesky.
// *noncvx (device) is raised when K is not positive definite.
static void band_refactor(BandDev &b, const DevMats &W, const double *rho, double sigma, int *noncvx, cudaStream_t st) {
  if (!b.ok) return;
  static std::once_flag once;
  std::call_once(once, [] {
    cudaFuncSetAttribute(k_band_factor, cudaFuncAttributeMaxDynamicSharedMemorySize, 5 * BAND_BMAX * (BAND_BMAX + 1) * 8);
  });
  k_band_init<<<grid_for((long long)b.npad * b.B), OSQP_CUDA_THREADS, 0, st>>>(b.new2old, b.B, b.npad, b.G, b.H);
  k_band_build<<<grid_for((long long)b.n * 32, 256), 256, 8 * 3 * b.B * sizeof(double), st>>>(
      view_of(W.M), view_of(W.A), rho, sigma, b.n, b.B, b.old2new, b.G, b.H);
  k_band_factor<<<1, BAND_THREADS, 5 * (size_t)b.B * (b.B + 1) * sizeof(double), st>>>(b.B, b.nl, b.G, b.H, b.JT, noncvx);
}
