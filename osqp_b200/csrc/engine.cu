// engine.cu -- the device-resident ADMM engine and the cuda_pcg plugin instance.
// C-ABI declared in include/osqp_cuda_pcg.h; reference lines cited per function.
#include "pcg.cuh"
#include "slack.cuh"
#include "dist.cuh"
#include "admm_fused.cuh"
#include <cmath>
#include <algorithm>
#include <mutex>
#include <pthread.h>
#include <time.h>

extern "C" int osqp_cuda_debug_set(int flags) {
  return cudaMemcpyToSymbol(g_spmv_dbg, &flags, sizeof(int)) != cudaSuccess;
}

// ===================================================================== launcher
// Counts launches and, when profiling is on, brackets each launch with CUDA
// events on the engine's own stream (bench.py's roofline numbers come from here).
struct Launcher {
  cudaStream_t st = 0;
  long long launches = 0;
  long long sell_launches = 0;   // SpMV launches over the sliced-ELL mirror
  long long tile_launches = 0;   // SpMV launches over a column-blocked mirror
  bool profiling = false;
  bool failed = false;           // a collective of the row-partitioned mode failed: surfaces at the next check
  struct Rec { int kid; cudaEvent_t a, b; bool own_a; };
  std::vector<Rec> recs;
  std::vector<cudaEvent_t> pool;
  double ms[OSQP_CUDA_K_COUNT] = {0};
  long long cnt[OSQP_CUDA_K_COUNT] = {0};
  cudaEvent_t cur_a = nullptr;
  int cur_kid = -1;
  // row-partitioned mode: SpMVs over `reduce_mat` (= M) are joined across the ranks
  DistCtx *dist = nullptr;
  double *dist_buf = nullptr;          // rows(M) x 4 doubles
  const CsrDev *reduce_mat = nullptr;

  cudaEvent_t get_event() {
    if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
  }
  void pre(int kid) {
    ++launches;
    if (profiling && kid >= 0) {
      cur_a = get_event(); cur_kid = kid;
      cudaEventRecord(cur_a, st);
    }
  }
  // sub-interval [start of the current launch group, now) under `kid2`: the tile kernel alone
  // inside a tiled SpMV (tile kernel + epilogue kernel are one logical kernel under `kid`)
  void mid(int kid2) {
    ++launches;
    if (profiling && cur_kid >= 0) {
      cudaEvent_t m = get_event();
      cudaEventRecord(m, st);
      recs.push_back({kid2, cur_a, m, false});
    }
  }
  void post() {
    if (profiling && cur_kid >= 0) {
      cudaEvent_t b = get_event();
      cudaEventRecord(b, st);
      recs.push_back({cur_kid, cur_a, b, true});
      cur_kid = -1;
      if (recs.size() >= 4096) collect();
    }
  }
  void collect() {
    if (recs.empty()) return;
    cudaStreamSynchronize(st);
    for (auto &r : recs) {
      float t = 0.f;
      cudaEventElapsedTime(&t, r.a, r.b);
      ms[r.kid] += t; cnt[r.kid] += 1;
      if (r.own_a) pool.push_back(r.a);
      pool.push_back(r.b);
    }
    recs.clear();
  }
  void destroy() {
    collect();
    for (auto e : pool) cudaEventDestroy(e);
    pool.clear();
  }
};

template <int NV, bool SPLIT, int MODE, class Epi>
static void launch_stream(Launcher &L, int kid, const CsrDev &C, int grid, Gather g0, Gather g1,
                          const Epi &epi, double *partials);

// device flag that turns a launch into a no-op (only the CG operator has one)
template <class E> static const int *epi_skip(const E &) { return nullptr; }
static const int *epi_skip(const EpiKp &e) { return &e.S->done; }

// Row-partitioned mode: partial row sums of M_g = [P_g | A_g'] -> allreduce(sum) -> epilogue.
template <int NV, bool SPLIT, int MODE, class Epi>
static void launch_reduced(Launcher &L, int kid, const CsrDev &C, int grid, Gather g0, Gather g1,
                           const Epi &epi, double *partials) {
  constexpr int NACC = NV * (SPLIT ? 2 : 1);
  DistCtx *d = L.dist;
  L.dist = nullptr;                      // the inner launch is a plain local SpMV
  EpiStoreSums<NACC> es{L.dist_buf, epi_skip(epi)};
  launch_stream<NV, SPLIT, MODE>(L, kid, C, grid, g0, g1, es, nullptr);
  L.dist = d;
  L.pre(OSQP_CUDA_K_ALLREDUCE);
  --L.launches;                          // a collective, not one of this library's kernels
  if (dist_allreduce(d, L.dist_buf, (size_t)C.rows * NACC, false, false, L.st)) L.failed = true;
  L.post();
  L.pre(OSQP_CUDA_K_APPLY_EPI);
  k_apply_epi<NACC, Epi><<<grid, OSQP_CUDA_THREADS, 0, L.st>>>(epi, L.dist_buf, C.rows, partials);
  L.post();
}

// A column-blocked mirror serves this launch (spmv_tile.cuh): plain products over one gathered
// vector, whitelisted epilogue, mirror current, not the matrix whose sums are joined across ranks.
static bool tile_path(const Launcher &L, const CsrDev &C) {
  return C.tile.ok && !C.tile.dirty && !(L.dist && &C == L.reduce_mat);
}
// per-CTA partial rows that a (NV 1, no split, MODE 0, whitelisted epilogue) launch over C produces
static int stream_parts(const Launcher &L, const CsrDev &C, int grid) {
  if (!tile_path(L, C)) return grid;
  return C.tile.rows_mode ? C.tile.ncta : tile_epi_grid(C.tile.nwin);
}
// L2 persistence for the partial sums of the PARTIAL-mode mirrors: the tile kernel streams > 100 MB
// through L2 while it writes them, and without a persisting window the combine kernel re-reads all
// of them from DRAM (ncu: 60 MB of DRAM reads, L2 hit rate 20 %).
static size_t tile_persist_bytes() {
  static std::once_flag once;
  static size_t window = 0;
  std::call_once(once, [] {
    // opt-in: measured on BASELINE config 2 it changes nothing for the combine kernel and costs the
    // other kernels their share of L2 (profiles/r2_bench_tile_v5_*.json)
    const char *env = getenv("OSQP_B200_TILE_PERSIST");
    if (!(env && env[0] == '1')) return;
    int dev = 0, max_persist = 0, max_window = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, dev) != cudaSuccess) { cudaGetLastError(); return; }
    const size_t want = std::min((size_t)max_persist, (size_t)64 << 20);
    if (want == 0 || cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) != cudaSuccess) { cudaGetLastError(); return; }
    window = std::min(want, (size_t)max_window);
  });
  return window;
}
template <class... KArgs, class... Args>
static void tile_launch_ex(void (*kern)(KArgs...), int grid, int block, size_t smem, cudaStream_t st,
                           const void *persist, size_t persist_bytes, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  const size_t window = tile_persist_bytes();
  if (persist && window) {
    at[0].id = cudaLaunchAttributeAccessPolicyWindow;
    at[0].val.accessPolicyWindow.base_ptr = const_cast<void *>(persist);
    at[0].val.accessPolicyWindow.num_bytes = std::min(persist_bytes, window);
    at[0].val.accessPolicyWindow.hitRatio = 1.0f;
    at[0].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    at[0].val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    cfg.attrs = at; cfg.numAttrs = 1;
  }
  cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

template <class Epi>
static bool launch_tiled(Launcher &L, int kid, const CsrDev &C, Gather g0, const Epi &epi, double *partials) {
  const TileDev &D = C.tile;
  // M = [P | A']: columns below n_lo gather from g0.lo, the others from g0.hi; A has one range, from g0.lo
  const double *xlo = D.nb_lo ? g0.lo : nullptr, *xhi = D.nb_lo ? g0.hi : g0.lo;
  if ((((size_t)xlo | (size_t)xhi) & 15) != 0) return false;     // TMA bulk copies need 16-byte aligned sources
  const TileView V = tile_view_of(D);
  const size_t smem = tile_smem_bytes(D);
  if (D.rows_mode) {
    static std::once_flag once;
    std::call_once(once, [] {
      cudaFuncSetAttribute(spmv_tile_rows_kernel<Epi>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE_SMEM_MAX);
    });
    L.pre(kid);
    ++L.tile_launches;
    spmv_tile_rows_kernel<Epi><<<V.ncta, TILE_THREADS, smem, L.st>>>(V, xlo, xhi, epi, partials);
    L.post();
    return true;
  }
  static std::once_flag once;
  std::call_once(once, [] {
    cudaFuncSetAttribute(spmv_tile_kernel<Epi>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE_SMEM_MAX);
  });
  const size_t pbytes = (size_t)D.nblocks * D.rows_pad * sizeof(double);
  L.pre(kid);
  ++L.tile_launches;
  tile_launch_ex(spmv_tile_kernel<Epi>, V.ncta, TILE_THREADS, smem, L.st, D.partial, pbytes, V, xlo, xhi, D.partial, epi);
  L.mid(D.nb_lo ? OSQP_CUDA_K_TILE_M : OSQP_CUDA_K_TILE_A);
  tile_launch_ex(k_tile_epi<Epi>, tile_epi_grid(V.nwin), TILE_WIN_PART, 0, L.st, D.partial, pbytes, epi, V,
                 (const double *)D.partial, partials);
  L.post();
  return true;
}

template <int NV, bool SPLIT, int MODE, class Epi>
static void launch_stream(Launcher &L, int kid, const CsrDev &C, int grid, Gather g0, Gather g1,
                          const Epi &epi, double *partials) {
  if (L.dist && &C == L.reduce_mat) {
    launch_reduced<NV, SPLIT, MODE>(L, kid, C, grid, g0, g1, epi, partials);
    return;
  }
  if constexpr (NV == 1 && !SPLIT && MODE == 0 && epi_tile_ok<Epi>::value) {
    if (tile_path(L, C) && launch_tiled(L, kid, C, g0, epi, partials)) return;
  }
  if (C.sell.ok && !C.sell.dirty && grid == C.sell.grid) {
    // sliced-ELL mirror, one contiguous stream per warp (spmv_sell.cuh)
    static std::once_flag sell_once;     // osqp_solve_batch worker threads come through here concurrently
    const size_t smem = sell_smem_bytes();
    std::call_once(sell_once, [smem] {
      if (smem)
        cudaFuncSetAttribute(spmv_sell_kernel<NV, SPLIT, MODE, Epi>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      else   // no shared memory on the row path: give L1 everything (gather lines in flight)
        cudaFuncSetAttribute(spmv_sell_kernel<NV, SPLIT, MODE, Epi>,
                             cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxL1);
    });
    const SellView V = sell_view_of(C);
    L.pre(kid);
    ++L.sell_launches;
    spmv_sell_kernel<NV, SPLIT, MODE, Epi>
        <<<C.sell.grid, OSQP_CUDA_THREADS, smem, L.st>>>(V, g0, g1, epi, partials);
    L.post();
    return;
  }
  static std::once_flag carve_once;
  // the kernel uses no shared memory: give L1 everything, part of the gathered vector lives there
  std::call_once(carve_once, [] {
    cudaFuncSetAttribute(spmv_direct_kernel<NV, SPLIT, MODE, Epi>,
                         cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxL1);
  });
  L.pre(kid);
  spmv_direct_kernel<NV, SPLIT, MODE, Epi>
      <<<grid, OSQP_CUDA_THREADS, 0, L.st>>>(view_of(C), g0, g1, epi, partials);
  L.post();
}

static int elementwise_grid(int len) {
  int g = (len + OSQP_CUDA_THREADS - 1) / OSQP_CUDA_THREADS;
  if (g < 1) g = 1;
  if (g > OSQP_CUDA_SMS * 4) g = OSQP_CUDA_SMS * 4;
  return g;
}

__global__ void pcg_set_cap_kernel(PcgScal *S, double cap) { S->tol_cap = cap; }
__global__ void pcg_reset_interval_kernel(PcgScal *S) { S->maxed = 0; S->max_iters = 0; }

// ================================================================ PCG driver
static int pcg_alloc(PcgDev &S, DevMats *W, cudaStream_t st) {
  S.W = W; S.n = W->n; S.m = W->m; S.st = st;
  if (dev_alloc(&S.Minv, S.n) || dev_alloc(&S.r, S.n) || dev_alloc(&S.p, S.n) ||
      dev_alloc(&S.Kp, S.n) || dev_alloc(&S.z, S.n) || dev_alloc(&S.w, S.m) ||
      dev_alloc(&S.partRZ[0], 2 * OSQP_MAX_PARTIAL_BLOCKS) ||
      dev_alloc(&S.partRZ[1], 2 * OSQP_MAX_PARTIAL_BLOCKS) ||
      dev_alloc(&S.partPKP, OSQP_MAX_PARTIAL_BLOCKS) || dev_alloc(&S.S, 1)) return 1;
  cudaMemsetAsync(S.S, 0, sizeof(PcgScal), st);
  pcg_set_cap_kernel<<<1, 1, 0, st>>>(S.S, DBL_MAX);
  cudaMemsetAsync(S.w, 0, sizeof(double) * (S.m ? S.m : 1), st);
  S.grid_n = elementwise_grid(S.n);
  S.gridA = spmv_grid(W->A, 4);
  S.gridM = spmv_grid(W->M, 4);
  return 0;
}
static void pcg_release(PcgDev &S) {
  dev_free(S.Minv); dev_free(S.r); dev_free(S.p); dev_free(S.Kp); dev_free(S.z); dev_free(S.w);
  dev_free(S.partRZ[0]); dev_free(S.partRZ[1]); dev_free(S.partPKP); dev_free(S.S);
}

// Minv <- 1 ./ (diag(P) + sigma + A.^2' rho)
static void pcg_jacobi(PcgDev &S, Launcher &L) {
  EpiJacobi ej{S.S, S.Minv, S.sigma};
  launch_stream<1, false, 1>(L, -1, S.W->M, S.gridM, Gather{nullptr, S.rho}, Gather{nullptr, nullptr},
                             ej, nullptr);
}

// Enqueue CG iterations [it0, it0+count) on x.  nparts0 = CTAs that produced partRZ[0].
static void pcg_enqueue(PcgDev &S, Launcher &L, double *x, int it0, int count, int nparts0,
                        double *zacc = nullptr) {
  // ADMM solves (zacc): the direction update is folded into K_b (EpiW::fuse), which then gathers z
  static const int fuse_env = [] { const char *e = getenv("OSQP_B200_PCG_FUSE_DIR"); return !(e && e[0] == '0'); }();
  // (a PARTIAL-mode mirror of A -- OSQP_B200_TILE_A=partial -- runs begin() in two kernels: not with the fold)
  const int fuse = zacc != nullptr && fuse_env && !(S.W->A.tile.ok && !S.W->A.tile.rows_mode);
  for (int it = it0; it < it0 + count; ++it) {
    EpiW ew{S.S, S.partRZ[it & 1], it == 0 ? nparts0 : S.grid_n, it, S.max_iter, S.rel_tol,
            S.abs_floor, S.use_cap, S.rho, S.w, zacc, fuse, S.n, S.Minv, S.r, S.p, 0.0, 0.0};
    launch_stream<1, false, 0>(L, OSQP_CUDA_K_SPMV_A, S.W->A, S.gridA, Gather{(fuse && it > 0) ? S.z : S.p, nullptr},
                               Gather{nullptr, nullptr}, ew, nullptr);
    EpiKp ek{S.S, S.p, S.Kp, S.sigma};
    launch_stream<1, false, 0>(L, OSQP_CUDA_K_SPMV_M, S.W->M, S.gridM, Gather{S.p, S.w},
                               Gather{nullptr, nullptr}, ek, S.partPKP);
    L.pre(OSQP_CUDA_K_PCG_UPDATE);
    pcg_update_kernel<<<S.grid_n, OSQP_CUDA_THREADS, 0, L.st>>>(
        S.S, S.partPKP, stream_parts(L, S.W->M, S.gridM), S.n, x, S.r, S.p, S.Kp, S.Minv, S.partRZ[(it + 1) & 1],
        fuse ? S.z : nullptr);
    L.post();
    if (fuse) continue;
    L.pre(OSQP_CUDA_K_PCG_DIR);
    pcg_dir_kernel<<<S.grid_n, OSQP_CUDA_THREADS, 0, L.st>>>(S.S, S.partRZ[(it + 1) & 1], S.grid_n,
                                                            S.n, S.p, S.r, S.Minv);
    L.post();
  }
}

// Host-synchronised solve of K x = b (strict plugin mode, polish, curvature probe).
// x holds the initial guess on entry unless x0_is_zero.  Returns CG iterations, <0 on error.
static int pcg_solve_sync(PcgDev &S, Launcher &L, const double *b, double *x, bool x0_is_zero,
                          double rel_tol, int max_iter, int *converged) {
  const double save_rel = S.rel_tol; const int save_max = S.max_iter, save_cap = S.use_cap;
  S.rel_tol = rel_tol; S.max_iter = max_iter; S.use_cap = 0;
  if (x0_is_zero) {
    cudaMemsetAsync(x, 0, sizeof(double) * S.n, L.st);
  } else {
    EpiWPlain ew{S.rho, S.w};
    launch_stream<1, false, 0>(L, -1, S.W->A, S.gridA, Gather{x, nullptr}, Gather{nullptr, nullptr},
                               ew, nullptr);
  }
  EpiInitGeneral ei{S.S, b, x, S.Minv, S.r, S.p, S.sigma, x0_is_zero ? 1 : 0};
  launch_stream<1, false, 0>(L, -1, S.W->M, S.gridM, Gather{x, S.w}, Gather{nullptr, nullptr}, ei,
                             S.partRZ[0]);
  PcgScal h;
  int it = 0, batch = 8;
  for (;;) {
    int todo = max_iter + 1 - it;      // the extra pass only runs the convergence test
    if (todo > batch) todo = batch;
    if (todo <= 0) break;
    pcg_enqueue(S, L, x, it, todo, stream_parts(L, S.W->M, S.gridM));
    it += todo;
    if (cudaMemcpyAsync(&h, S.S, sizeof(PcgScal), cudaMemcpyDeviceToHost, L.st) != cudaSuccess ||
        cudaStreamSynchronize(L.st) != cudaSuccess) {
      S.rel_tol = save_rel; S.max_iter = save_max; S.use_cap = save_cap;
      return -1;
    }
    if (h.done) break;
    if (batch < 64) batch *= 2;
  }
  L.pre(-1);
  pcg_close_kernel<<<1, OSQP_CUDA_THREADS, 0, L.st>>>(S.S, S.partRZ[it & 1],
                                                      it == 0 ? stream_parts(L, S.W->M, S.gridM) : S.grid_n);
  L.post();
  cudaMemcpyAsync(&h, S.S, sizeof(PcgScal), cudaMemcpyDeviceToHost, L.st);
  cudaError_t e = cudaStreamSynchronize(L.st);
  S.rel_tol = save_rel; S.max_iter = save_max; S.use_cap = save_cap;
  if (e != cudaSuccess) {
    fprintf(stderr, "osqp_cuda_pcg: PCG solve failed: %s\n", cudaGetErrorString(e));
    return -1;
  }
  if (converged) *converged = h.conv;
  return h.iters;
}

// ==================================================================== engine
#define RESM_NSUM 1
#define RESM_NMAX 12
#define RESN_NSUM 3
#define RESN_NMAX 14
#define OSQP_PCG_EXACT_N 128                  // n at or below which the CG runs to full accuracy
#define OSQP_BIG (OSQP_INFTY * MIN_SCALING)   // "infinite bound" threshold, reference auxil.c:82-83

struct osqp_cuda_engine {
  DevMats W;
  PcgDev pcg;
  Launcher L;
  cudaStream_t st = 0;
  int n = 0, m = 0;
  bool linsys_ready = false;
  cuda_pcg_solver *solver = nullptr;   // plugin view of this engine's linear system
  // data (scaled) and raw copies
  double *q = nullptr, *l = nullptr, *u = nullptr;
  double *q_raw = nullptr, *l_raw = nullptr, *u_raw = nullptr;
  // scaling
  double *D = nullptr, *Dinv = nullptr, *E = nullptr, *Einv = nullptr;
  double *Dtmp = nullptr, *Etmp = nullptr;
  double *cs = nullptr;               // device scalars: [0] c, [1] cinv, [2] c_temp
  double c_host = 1.0;
  int scaling_iters = 0;
  // rho
  double rho = 0.1, sigma = 1e-6;
  double *rho_vec = nullptr;           // m
  int *ctype = nullptr;                // m: -1 loose, 0 inequality, 1 equality
  int *flag = nullptr;                 // device int scratch
  // iterates
  double *x = nullptr, *xt = nullptr, *dx = nullptr;                 // n
  double *z = nullptr, *y = nullptr, *Ax = nullptr, *v = nullptr;    // m
  double *dy = nullptr, *dyp = nullptr;                              // m
  double *zacc = nullptr;              // m: A (x~ - x) accumulated inside the CG (EpiW)
  bool v_dirty = true;
  // polish
  double *pol_x = nullptr, *pol_z = nullptr, *pol_y = nullptr, *pol_rho = nullptr;
  double *pol_b = nullptr, *pol_rhs = nullptr, *pol_nu = nullptr, *pol_t = nullptr;
  int *pol_act = nullptr;
  // reductions
  double *partM = nullptr, *partN = nullptr;
  double *res_dev = nullptr;           // osqp_cuda_resid + pcg interval stats
  double *res_host = nullptr;          // pinned
  // PCG launch policy
  int pcg_launch = 6;                 // CG iterations enqueued per ADMM iteration
  int pcg_probe = 6;                  // ADMM iterations still to run in probe mode
  double pcg_rel = 0.1, pcg_abs = 1e-30, pcg_lambda = 0.05;
  int pcg_max = 500;
  OSQPB200Stats stats = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  double bytes_per_launch[OSQP_CUDA_K_COUNT] = {0};
  cudaEvent_t ev_a = nullptr, ev_b = nullptr;   // osqp_cuda_engine_timer_*
  // One ADMM iteration (rhs, L CG iterations, ADMM update: 4L + 2 launches with fixed
  // arguments) captured once per configuration and replayed as a CUDA graph.
  struct IterGraph {
    int L, store, maxit, use_cap, gridA, gridM;
    double alpha, rel, absf;
    long long launches;
    cudaGraphExec_t exec;
  };
  std::vector<IterGraph> graphs;
  bool graphs_on = true;
  bool fused = false;                 // small problem: whole ADMM iterations in one single-CTA kernel
  BandDev band;                       // fused problems with a banded K: exact block-tridiagonal preconditioner
  SlackDev slack;                     // slack variables eliminated from the CG's system (slack.cuh)
  DistCtx *dist = nullptr;            // row-partitioned mode (null: single GPU)
  double *dist_buf = nullptr, *partMc = nullptr;
  bool warmed = false;                // one uncaptured iteration has run (function attributes set)
};

// ------------------------------------------------------------ ADMM epilogues
// K_a (reference compute_rhs, src/auxil.c:161-175, in reduced form with the CG
// started from the current iterate):  r0 = rhs - K x = -(q + P x + A' v),
// v = y + rho .* (A x - z);  p = Minv r0;  x~ <- x
struct EpiRhs {
  static constexpr int NSUM = 1, NMAX = 1;
  PcgScal *S;
  const double *q, *x, *Minv;
  double *r, *p, *xt;
  __device__ bool begin(double *) {
    if (blockIdx.x == 0 && threadIdx.x == 0) { S->done = 0; S->conv = 0; S->iters = 0; S->applied = 0; }
    return true;
  }
  struct Pre { double q, x, Minv; };
  __device__ __forceinline__ Pre load(int j) const { return Pre{q[j], x[j], Minv[j]}; }
  __device__ void row(int j, const double *s, const Pre &pre, double *asum, double *amax) const {
    const double rj = pre.Minv == 0.0 ? 0.0 : -(pre.q + s[0]);    // Minv == 0: eliminated slack (slack.cuh)
    const double zj = pre.Minv * rj;
    r[j] = rj; p[j] = zj; xt[j] = pre.x;
    asum[0] += rj * zj;
    amax[0] = fmax(amax[0], fabs(rj));
  }
};

// K_f = K7 (reference update_x / update_z + project / update_y, src/auxil.c:185-225,
// src/proj.c:4-14) without an SpMV: z~ = A x~ = Ax + zacc + alpha_last * (w ./ rho)   (the last
// CG step's A p is still in w; see EpiW), then x, z, y and the bookkeeping that makes the next
// K_a free of extra SpMVs (A x by recurrence, v).  One pass over the m-vectors instead of a pass over A.
__global__ void __launch_bounds__(OSQP_CUDA_THREADS)
k_admm_update(PcgScal *S, const double *part_last, int nparts_last, int n, int m, double alpha, int store,
              double *x, const double *xt, double *dx, const double *rho, const double *l, const double *u,
              double *z, double *y, double *Ax, double *v, double *dy, const double *w, double *zacc,
              SlackView sk) {
  __shared__ double scratch[32];
  // read before block 0 closes the solve (pcg_close does not touch these fields)
  const int applied = S->applied, iters = S->iters;
  const double al = S->alpha_last;
  const bool pending = applied >= 1 && applied == iters;   // the last w's step is not in zacc yet
  if (blockIdx.x == 0) pcg_close(S, part_last, nparts_last, scratch);
  const int stride = gridDim.x * blockDim.x, t0 = blockIdx.x * blockDim.x + threadIdx.x;
  // With slack elimination (sk.ny > 0) the CG leaves x~ = x on the slacks; the thread of a slack's
  // partner row computes its step below and updates x_j and dx_j itself, this loop skips them.
  for (int j = t0; j < n; j += stride) {
    if (sk.ny && sk.minv[j] == 0.0) continue;
    const double xo = x[j];
    const double xn = alpha * xt[j] + (1.0 - alpha) * xo;
    x[j] = xn;
    if (store) dx[j] = xn - xo;
  }
  const double *rho_w = sk.ny ? sk.rho_cg : rho;      // the rho the CG scaled w with
  for (int i = t0; i < m; i += stride) {
    const double rh = rho[i], axo = Ax[i];
    double zt = zacc[i];                              // A_X (x~ - x)_X so far
    if (pending) zt += al * (w[i] / rho_w[i]);
    if (sk.ny) {
      const int t = sk.slot[i];
      if (t >= 0) {
        const int j = sk.var[t];
        const double aj = sk.a[t];
        const double dj = (sk.r0[t] - aj * rh * zt) / sk.D[t];
        zt += aj * dj;
        const double xo = x[j];
        x[j] = xo + alpha * dj;
        if (store) dx[j] = alpha * dj;
      }
    }
    zt += axo;
    zacc[i] = 0.0;
    const double zo = z[i], yo = y[i];
    const double rinv = 1.0 / rh;
    const double zr = alpha * zt + (1.0 - alpha) * zo;
    const double zn = fmin(fmax(zr + rinv * yo, l[i]), u[i]);
    const double d = rh * (zr - zn);
    const double yn = yo + d;
    const double axn = alpha * zt + (1.0 - alpha) * axo;
    z[i] = zn; y[i] = yn; Ax[i] = axn;
    v[i] = yn + rh * (axn - zn);
    if (store) dy[i] = d;
  }
}

// Ax <- A x, v <- y + rho .* (Ax - z)    (after warm start / rho change / polish)
struct EpiAxV {
  static constexpr int NSUM = 0, NMAX = 0;
  const double *rho, *z, *y;
  double *Ax, *v;
  __device__ bool begin(double *) { return true; }
  struct Pre { double rho, z, y; };
  __device__ __forceinline__ Pre load(int i) const { return Pre{rho[i], z[i], y[i]}; }
  __device__ void row(int i, const double *s, const Pre &pre, double *, double *) const {
    Ax[i] = s[0];
    v[i] = pre.y + pre.rho * (s[0] - pre.z);
  }
};
// z <- A x (warm start, reference src/osqp.c:962)
struct EpiStore {
  static constexpr int NSUM = 0, NMAX = 0;
  double *out;
  __device__ bool begin(double *) { return true; }
  EPI_NO_PREFETCH
  __device__ void row(int i, const double *s, const Pre &, double *, double *) const { out[i] = s[0]; }
};

template <> struct epi_tile_ok<EpiRhs> { static constexpr bool value = true; };
template <> struct epi_tile_ok<EpiAxV> { static constexpr bool value = true; };
template <> struct epi_tile_ok<EpiStore> { static constexpr bool value = true; };

// K9/K10, m side (reference src/auxil.c:240-285, 361-398, 475-497):
// gathers x and dx; sums: [ineq_lhs]; maxes: see osqp_cuda_engine_residuals
struct EpiResM {
  static constexpr int NSUM = RESM_NSUM, NMAX = RESM_NMAX;
  const double *z, *y, *rho, *l, *u, *E, *Einv, *dy;
  double *Ax, *v, *dyp;
  int write_back;
  __device__ bool begin(double *) { return true; }
  EPI_NO_PREFETCH
  __device__ void row(int i, const double *s, const Pre &, double *asum, double *amax) const {
    const double ax = s[0], adx = s[1], zi = z[i], ei = E[i], einv = Einv[i];
    const double li = l[i], ui = u[i];
    const double pr = ax - zi;
    amax[0] = fmax(amax[0], fabs(pr));
    amax[1] = fmax(amax[1], fabs(einv * pr));
    amax[2] = fmax(amax[2], fabs(zi));
    amax[3] = fmax(amax[3], fabs(einv * zi));
    amax[4] = fmax(amax[4], fabs(ax));
    amax[5] = fmax(amax[5], fabs(einv * ax));
    if (write_back) {
      Ax[i] = ax;
      v[i] = y[i] + rho[i] * pr;
      // projection of delta_y on the polar of the recession cone of [l,u]
      double d = dy[i];
      if (ui > OSQP_BIG) {
        d = (li < -OSQP_BIG) ? 0.0 : fmin(d, 0.0);
      } else if (li < -OSQP_BIG) {
        d = fmax(d, 0.0);
      }
      dyp[i] = d;
      amax[6] = fmax(amax[6], fabs(d));
      amax[7] = fmax(amax[7], fabs(ei * d));
      asum[0] += ui * fmax(d, 0.0) + li * fmin(d, 0.0);
      if (ui < OSQP_BIG) {
        amax[8] = fmax(amax[8], adx);
        amax[9] = fmax(amax[9], einv * adx);
      }
      if (li > -OSQP_BIG) {
        amax[10] = fmax(amax[10], -adx);
        amax[11] = fmax(amax[11], -(einv * adx));
      }
    }
  }
};

// K9/K10, n side (reference src/auxil.c:287-359, 227-238, 400-424, 426-470):
// gathers [x ; y] and [dx ; dyp]; split sums: Px, A'y, P dx, A' dyp
struct EpiResN {
  static constexpr int NSUM = RESN_NSUM, NMAX = RESN_NMAX;
  const double *q, *x, *dx, *D, *Dinv;
  __device__ bool begin(double *) { return true; }
  EPI_NO_PREFETCH
  __device__ void row(int j, const double *s, const Pre &, double *asum, double *amax) const {
    const double Px = s[0], Aty = s[1], Pdx = s[2], Atdy = s[3];
    const double qj = q[j], xj = x[j], dxj = dx[j], dinv = Dinv[j];
    const double d = (qj + Px) + Aty;
    amax[0] = fmax(amax[0], fabs(d));
    amax[1] = fmax(amax[1], fabs(dinv * d));
    amax[2] = fmax(amax[2], fabs(qj));
    amax[3] = fmax(amax[3], fabs(dinv * qj));
    amax[4] = fmax(amax[4], fabs(Aty));
    amax[5] = fmax(amax[5], fabs(dinv * Aty));
    amax[6] = fmax(amax[6], fabs(Px));
    amax[7] = fmax(amax[7], fabs(dinv * Px));
    amax[8] = fmax(amax[8], fabs(Atdy));
    amax[9] = fmax(amax[9], fabs(dinv * Atdy));
    amax[10] = fmax(amax[10], fabs(dxj));
    amax[11] = fmax(amax[11], fabs(D[j] * dxj));
    amax[12] = fmax(amax[12], fabs(Pdx));
    amax[13] = fmax(amax[13], fabs(dinv * Pdx));
    asum[0] += xj * Px;
    asum[1] += qj * xj;
    asum[2] += qj * dxj;
  }
};

// ------------------------------------------------------- elementwise kernels
#define EW_LOOP(i, len) \
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < (len); i += gridDim.x * blockDim.x)

__global__ void k_fill(double *a, double v, int len) { EW_LOOP(i, len) a[i] = v; }
__global__ void k_mul(double *out, const double *a, const double *b, int len) {
  EW_LOOP(i, len) out[i] = a[i] * b[i];
}
__global__ void k_recip(double *out, const double *a, int len) { EW_LOOP(i, len) out[i] = 1.0 / a[i]; }
// out = s * a .* b  with s read from device memory (s_dev may be null -> 1)
__global__ void k_mul_scal(double *out, const double *a, const double *b, const double *s_dev, int len) {
  const double s = s_dev ? *s_dev : 1.0;
  EW_LOOP(i, len) out[i] = (a[i] * b[i]) * s;
}
// Ruiz step on a norm vector (reference src/scaling.c:7-14,82-101):
// t <- 1/sqrt(limit(t)); cum <- cum .* t; optional q <- q .* t
__global__ void k_ruiz_vec(double *t, double *cum, double *q, int len) {
  EW_LOOP(i, len) {
    double a = t[i];
    a = a < MIN_SCALING ? 1.0 : a;
    a = a > MAX_SCALING ? MAX_SCALING : a;
    a = 1.0 / sqrt(a);
    t[i] = a;
    cum[i] *= a;
    if (q) q[i] *= a;
  }
}
// Cost normalisation scalar (reference src/scaling.c:107-141), one CTA:
// c_temp = 1 / limit(max(mean(colnorm P), limit(||q||inf)));  cs[0] *= c_temp; cs[2] = c_temp
__global__ void __launch_bounds__(1024) k_cost_scale(const double *pnorm, const double *q, int n, double *cs) {
  __shared__ double ssum[32], smax[32];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) { a += pnorm[i]; b = fmax(b, fabs(q[i])); }
  a = warp_sum(a); b = warp_max(b);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { ssum[wid] = a; smax[wid] = b; }
  __syncthreads();
  if (wid == 0) {
    a = lane < (blockDim.x >> 5) ? ssum[lane] : 0.0;
    b = lane < (blockDim.x >> 5) ? smax[lane] : 0.0;
    a = warp_sum(a); b = warp_max(b);
    if (lane == 0) {
      double c_temp = a / n;
      double nq = b;
      nq = nq < MIN_SCALING ? 1.0 : nq; nq = nq > MAX_SCALING ? MAX_SCALING : nq;
      c_temp = fmax(c_temp, nq);
      c_temp = c_temp < MIN_SCALING ? 1.0 : c_temp; c_temp = c_temp > MAX_SCALING ? MAX_SCALING : c_temp;
      c_temp = 1.0 / c_temp;
      cs[2] = c_temp;
      cs[0] *= c_temp;
      cs[1] = 1.0 / cs[0];
    }
  }
}
__global__ void k_scale_by_dev(double *a, const double *s_dev, int len) {
  const double s = *s_dev;
  EW_LOOP(i, len) a[i] *= s;
}
// Constraint types and rho_vec (reference src/auxil.c:76-142).  retype == 0: write
// everything; retype == 1: only rows whose type changed, and raise *changed.
__global__ void k_rho_types(const double *l, const double *u, double rho, int m, int *ctype,
                            double *rho_vec, int retype, int *changed) {
  EW_LOOP(i, m) {
    int t;
    double r;
    if (l[i] < -OSQP_BIG && u[i] > OSQP_BIG) { t = -1; r = RHO_MIN; }
    else if (u[i] - l[i] < RHO_TOL) { t = 1; r = RHO_EQ_OVER_RHO_INEQ * rho; }
    else { t = 0; r = rho; }
    if (!retype) { ctype[i] = t; rho_vec[i] = r; }
    else if (ctype[i] != t) { ctype[i] = t; rho_vec[i] = r; *changed = 1; }
  }
}
// reference src/osqp.c:1304-1316: loose rows keep RHO_MIN
__global__ void k_rho_update(const int *ctype, double rho, int m, double *rho_vec) {
  EW_LOOP(i, m) {
    if (ctype[i] == 0) rho_vec[i] = rho;
    else if (ctype[i] == 1) rho_vec[i] = RHO_EQ_OVER_RHO_INEQ * rho;
  }
}
// x-update alone (m == 0 has no row kernel to fuse it into)
__global__ void k_axpby_store(double *x, const double *xt, double *dx, double alpha, int store, int n) {
  EW_LOOP(j, n) {
    const double xo = x[j];
    const double xn = alpha * xt[j] + (1.0 - alpha) * xo;
    x[j] = xn;
    if (store) dx[j] = xn - xo;
  }
}
// Final stage of the residual reductions: one CTA, fixed order.
// out[0..] = sums then maxes of partM, then sums then maxes of partN, then PCG stats.
__global__ void __launch_bounds__(OSQP_CUDA_THREADS)
k_res_final(const double *partM, int nM, const double *partN, int nN, PcgScal *S, double *out,
            double lambda, int set_cap) {
  __shared__ double scratch[32];
  const int KM = RESM_NSUM + RESM_NMAX, KN = RESN_NSUM + RESN_NMAX;
  for (int k = 0; k < KM; ++k) {
    const bool is_max = k >= RESM_NSUM;
    double v = is_max ? -DBL_MAX : 0.0;
    for (int b = threadIdx.x; b < nM; b += blockDim.x) {
      const double t = partM[(size_t)b * KM + k];
      v = is_max ? fmax(v, t) : v + t;
    }
    v = is_max ? block_reduce<true>(v, scratch) : block_reduce<false>(v, scratch);
    if (threadIdx.x == 0) out[k] = v;
  }
  for (int k = 0; k < KN; ++k) {
    const bool is_max = k >= RESN_NSUM;
    double v = is_max ? -DBL_MAX : 0.0;
    for (int b = threadIdx.x; b < nN; b += blockDim.x) {
      const double t = partN[(size_t)b * KN + k];
      v = is_max ? fmax(v, t) : v + t;
    }
    v = is_max ? block_reduce<true>(v, scratch) : block_reduce<false>(v, scratch);
    if (threadIdx.x == 0) out[KM + k] = v;
  }
  if (threadIdx.x == 0) {
    out[KM + KN + 0] = (double)S->total_iters;
    out[KM + KN + 1] = (double)S->solves;
    out[KM + KN + 2] = (double)S->maxed;
    out[KM + KN + 3] = (double)S->max_iters;
    out[KM + KN + 4] = (double)S->noncvx;
    out[KM + KN + 5] = (double)S->breakdowns;
    S->maxed = 0;
    S->max_iters = 0;
    if (set_cap) {
      // CG target for the coming ADMM iterations: lambda * sqrt(pri_res * dua_res) in the scaled
      // space (the rule of the GPU OSQP paper cited at reference docs/citing/index.rst:45-57)
      const double pri = nM > 0 ? out[RESM_NSUM + 0] : 0.0, dua = out[KM + RESN_NSUM + 0];
      double cap = (pri > 0.0) ? lambda * sqrt(pri * dua) : lambda * dua;
      if (!(lambda > 0.0) || !(cap == cap)) cap = DBL_MAX;
      S->tol_cap = cap;
    }
  }
}
#define RES_OUT_LEN (RESM_NSUM + RESM_NMAX + RESN_NSUM + RESN_NMAX + 8)

// ============================================================ C-ABI: engine
extern "C" void free_linsys_solver_cuda_pcg(cuda_pcg_solver *self);
#define EW(len) elementwise_grid(len), OSQP_CUDA_THREADS, 0, e->st
static int g_device = -1;

extern "C" c_int osqp_cuda_set_device(c_int device) {
  g_device = (int)device;
  return cudaSetDevice(g_device) != cudaSuccess;
}
static int bind_device() {
  const bool first = g_device < 0;
  if (first) {
    // Concurrent solves (the threaded path of osqp_solve_batch) want more than the default 8 hardware
    // queues.  Only before this library's first CUDA call, never over a value the application set, and
    // not at all with OSQP_B200_KEEP_ENV=1 (the one process-wide side effect of loading this backend).
    if (!getenv("OSQP_B200_KEEP_ENV")) setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
    const char *env = getenv("OSQP_CUDA_DEVICE");
    g_device = env ? atoi(env) : 0;
  }
  cudaError_t err = cudaSetDevice(g_device);
  if (err != cudaSuccess) {
    fprintf(stderr, "osqp_cuda_pcg: no usable CUDA device %d: %s (this backend has no CPU fallback)\n",
            g_device, cudaGetErrorString(err));
    return 1;
  }
  if (first) {
    // OSQP_CUDA_BLOCKING_SYNC=1: host threads sleep in stream synchronisation instead of
    // spinning (many osqp_solve_batch threads on few cores); off by default because every
    // termination check of a single solve would pay the wake-up latency
    const char *bs = getenv("OSQP_CUDA_BLOCKING_SYNC");
    if (bs && bs[0] == '1' && cudaSetDeviceFlags(cudaDeviceScheduleBlockingSync) != cudaSuccess)
      cudaGetLastError();
  }
  return 0;
}
// ---- row-partitioned mode bootstrap (include/osqp_cuda_pcg.h)
extern "C" c_int osqp_cuda_nccl_unique_id(unsigned char *id128) {
  if (!id128 || nccl_load()) return 1;
  ncclUniqueId id;
  if (g_nccl.GetUniqueId(&id) != ncclSuccess) return 1;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  memcpy(id128, &id, 128);
  return 0;
}
extern "C" c_int osqp_cuda_set_partition(c_int rank, c_int world, const unsigned char *id128) {
  if (g_dist.comm) { g_nccl.CommDestroy(g_dist.comm); g_dist = DistCtx(); }
  if (world <= 1) return 0;
  if (!id128 || rank < 0 || rank >= world || nccl_load() || bind_device()) return 1;
  ncclUniqueId id;
  memcpy(&id, id128, 128);
  ncclComm_t comm = nullptr;
  const ncclResult_t r = g_nccl.CommInitRank(&comm, (int)world, id, (int)rank);
  if (r != ncclSuccess) {
    fprintf(stderr, "osqp_cuda_pcg: ncclCommInitRank failed: %s\n", g_nccl.GetErrorString(r));
    return 1;
  }
  g_dist.comm = comm; g_dist.rank = (int)rank; g_dist.world = (int)world; g_dist.allreduces = 0;
  return 0;
}
// world of ONE with the partitioned code path switched on: every [P | A'] product stores its row
// sums, "joins" them (identity) and runs its epilogue in the separate kernel, exactly as with
// world > 1.  For single-GPU tests of that path (the collective itself needs >= 2 GPUs).
extern "C" c_int osqp_cuda_set_partition_loopback(c_int on) {
  if (g_dist.comm) { g_nccl.CommDestroy(g_dist.comm); }
  g_dist = DistCtx();
  g_dist.loopback = on != 0;
  return 0;
}
extern "C" c_int osqp_cuda_get_partition(c_int *rank, c_int *world, c_int *allreduces) {
  if (rank) *rank = g_dist.rank;
  if (world) *world = g_dist.world;
  if (allreduces) *allreduces = g_dist.allreduces;
  return 0;
}

extern "C" c_int osqp_cuda_probe(void) {
  if (bind_device()) return 1;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, g_device) != cudaSuccess) return 1;
  if (prop.major != 10) {
    fprintf(stderr, "osqp_cuda_pcg: built for sm_100a only, found sm_%d%d\n", prop.major, prop.minor);
    return 1;
  }
  return 0;
}

extern "C" void osqp_cuda_engine_free(osqp_cuda_engine *e) {
  if (!e) return;
  if (e->st) cudaStreamSynchronize(e->st);
  for (auto &g : e->graphs) cudaGraphExecDestroy(g.exec);
  e->graphs.clear();
  e->L.destroy();
  if (e->solver) free_linsys_solver_cuda_pcg(e->solver);   // the bound view, if still alive
  pcg_release(e->pcg);
  band_free(e->band);
  mats_free(e->W);
  dev_free(e->q); dev_free(e->l); dev_free(e->u); dev_free(e->q_raw); dev_free(e->l_raw);
  dev_free(e->u_raw); dev_free(e->D); dev_free(e->Dinv); dev_free(e->E); dev_free(e->Einv);
  dev_free(e->Dtmp); dev_free(e->Etmp); dev_free(e->cs); dev_free(e->rho_vec); dev_free(e->ctype);
  dev_free(e->flag); dev_free(e->x); dev_free(e->xt); dev_free(e->dx); dev_free(e->z);
  dev_free(e->y); dev_free(e->Ax); dev_free(e->v); dev_free(e->dy); dev_free(e->dyp); dev_free(e->zacc);
  slack_free(e->slack);
  dev_free(e->pol_x); dev_free(e->pol_z); dev_free(e->pol_y); dev_free(e->pol_rho);
  dev_free(e->pol_b); dev_free(e->pol_rhs); dev_free(e->pol_nu); dev_free(e->pol_t);
  dev_free(e->pol_act); dev_free(e->partM); dev_free(e->partN); dev_free(e->res_dev);
  dev_free(e->dist_buf); dev_free(e->partMc);
  if (e->res_host) cudaFreeHost(e->res_host);
  if (e->ev_a) { cudaEventDestroy(e->ev_a); cudaEventDestroy(e->ev_b); }
  if (e->st) cudaStreamDestroy(e->st);
  delete e;
}

static void engine_reset_scaling(osqp_cuda_engine *e) {
  const int n = e->n, m = e->m;
  k_fill<<<EW(n)>>>(e->D, 1.0, n);
  k_fill<<<EW(n)>>>(e->Dinv, 1.0, n);
  if (m) { k_fill<<<EW(m)>>>(e->E, 1.0, m); k_fill<<<EW(m)>>>(e->Einv, 1.0, m); }
  const double one[3] = {1.0, 1.0, 1.0};
  cudaMemcpyAsync(e->cs, one, sizeof(one), cudaMemcpyHostToDevice, e->st);
  cudaStreamSynchronize(e->st);
  e->c_host = 1.0;
}

extern "C" c_int osqp_cuda_engine_create(osqp_cuda_engine **ep, c_int n, c_int m, const csc *P,
                                         const csc *A, const c_float *q, const c_float *l,
                                         const c_float *u) {
  *ep = nullptr;
  if (bind_device()) return 1;
  osqp_cuda_engine *e = new osqp_cuda_engine();
  e->n = (int)n; e->m = (int)m;
  if (cudaStreamCreateWithFlags(&e->st, cudaStreamNonBlocking) != cudaSuccess) {
    fprintf(stderr, "osqp_cuda_pcg: cannot create a CUDA stream: %s\n",
            cudaGetErrorString(cudaGetLastError()));
    delete e;
    return 1;
  }
  e->L.st = e->st;
  int bad = mats_build(e->W, P, A, e->st);
  const size_t N = (size_t)n, Mz = (size_t)m;
  bad = bad || dev_alloc(&e->q, N) || dev_alloc(&e->q_raw, N) || dev_alloc(&e->l, Mz) ||
        dev_alloc(&e->u, Mz) || dev_alloc(&e->l_raw, Mz) || dev_alloc(&e->u_raw, Mz) ||
        dev_alloc(&e->D, N) || dev_alloc(&e->Dinv, N) || dev_alloc(&e->E, Mz) ||
        dev_alloc(&e->Einv, Mz) || dev_alloc(&e->Dtmp, N) || dev_alloc(&e->Etmp, Mz) ||
        dev_alloc(&e->cs, 4) || dev_alloc(&e->rho_vec, Mz) || dev_alloc(&e->ctype, Mz) ||
        dev_alloc(&e->flag, 4) || dev_alloc(&e->x, N) || dev_alloc(&e->xt, N) ||
        dev_alloc(&e->dx, N) || dev_alloc(&e->z, Mz) || dev_alloc(&e->y, Mz) ||
        dev_alloc(&e->Ax, Mz) || dev_alloc(&e->v, Mz) || dev_alloc(&e->dy, Mz) ||
        dev_alloc(&e->dyp, Mz) || dev_alloc(&e->zacc, Mz) ||
        dev_alloc(&e->partM, (size_t)OSQP_MAX_PARTIAL_BLOCKS * (RESM_NSUM + RESM_NMAX)) ||
        dev_alloc(&e->partN, (size_t)OSQP_MAX_PARTIAL_BLOCKS * (RESN_NSUM + RESN_NMAX)) ||
        dev_alloc(&e->res_dev, RES_OUT_LEN);
  if (!bad) bad = cudaMallocHost((void **)&e->res_host, sizeof(double) * RES_OUT_LEN) != cudaSuccess;
  if (!bad) bad = pcg_alloc(e->pcg, &e->W, e->st);
  if (!bad && (g_dist.world > 1 || g_dist.loopback)) {
    e->dist = &g_dist;
    bad = dev_alloc(&e->dist_buf, 4 * (N ? N : 1)) || dev_alloc(&e->partMc, RESM_NSUM + RESM_NMAX);
    e->L.dist = e->dist; e->L.dist_buf = e->dist_buf; e->L.reduce_mat = &e->W.M;
    // the NCCL calls are captured into the iteration graphs like the kernels (OSQP_B200_DIST_GRAPHS=0: plain launches)
    { const char *g = getenv("OSQP_B200_DIST_GRAPHS"); e->graphs_on = !(g && g[0] == '0'); }
  }
  if (bad) { osqp_cuda_engine_free(e); return 1; }
  {
    // small problems (all iterates + CG vectors fit one SM's shared memory): fused single-CTA path
    const char *fz = getenv("OSQP_B200_FUSED");
    e->fused = !e->dist && !(fz && fz[0] == '0') && admm_fused_smem((int)n, (int)m) <= ADMM_FUSED_MAX_SMEM;
    if (e->fused &&
        (cudaFuncSetAttribute(admm_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              ADMM_FUSED_MAX_SMEM) != cudaSuccess ||
         cudaFuncSetAttribute(admm_fused_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              ADMM_FUSED_MAX_SMEM) != cudaSuccess)) {
      cudaGetLastError();
      e->fused = false;
    }
    // banded K (MPC-type problems): exact block-tridiagonal preconditioner for the fused kernel (band.cuh)
    if (e->fused) {
      if (band_setup(e->band, P, A, e->st)) { osqp_cuda_engine_free(e); return 1; }
      if (e->band.ok && admm_fused_smem((int)n, (int)m) + band_apply_smem(e->band) > ADMM_FUSED_MAX_SMEM) band_free(e->band);
    }
  }
  // slack variables (Lasso-type problems): eliminated exactly from the CG's system (slack.cuh); the
  // multi-kernel path only
  if (!e->fused && !e->dist && slack_setup(e->slack, P, A, e->W, e->st)) { osqp_cuda_engine_free(e); return 1; }
  e->pcg.rho = e->rho_vec;
  cudaMemcpyAsync(e->q_raw, q, sizeof(double) * N, cudaMemcpyHostToDevice, e->st);
  if (m) {
    cudaMemcpyAsync(e->l_raw, l, sizeof(double) * Mz, cudaMemcpyHostToDevice, e->st);
    cudaMemcpyAsync(e->u_raw, u, sizeof(double) * Mz, cudaMemcpyHostToDevice, e->st);
  }
  cudaMemcpyAsync(e->q, e->q_raw, sizeof(double) * N, cudaMemcpyDeviceToDevice, e->st);
  if (m) {
    cudaMemcpyAsync(e->l, e->l_raw, sizeof(double) * Mz, cudaMemcpyDeviceToDevice, e->st);
    cudaMemcpyAsync(e->u, e->u_raw, sizeof(double) * Mz, cudaMemcpyDeviceToDevice, e->st);
  }
  double *zero_n[] = {e->x, e->xt, e->dx};
  for (double *p : zero_n) cudaMemsetAsync(p, 0, sizeof(double) * N, e->st);
  double *zero_m[] = {e->z, e->y, e->Ax, e->v, e->dy, e->dyp, e->zacc};
  for (double *p : zero_m) cudaMemsetAsync(p, 0, sizeof(double) * (Mz ? Mz : 1), e->st);
  engine_reset_scaling(e);
  if (cudaStreamSynchronize(e->st) != cudaSuccess || cudaGetLastError() != cudaSuccess) {
    fprintf(stderr, "osqp_cuda_pcg: engine creation failed on the device\n");
    osqp_cuda_engine_free(e);
    return 1;
  }
  // algorithmic bytes per launch of the stream kernels (12 B/nnz + 4 B/row ptr
  // + gathered and written vectors counted once)
  const double bA = 12.0 * e->W.A.nnz + 4.0 * (m + 1) + 8.0 * n + 8.0 * m;
  const double bM = 12.0 * e->W.M.nnz + 4.0 * (n + 1) + 8.0 * (n + m) + 8.0 * n;
  e->bytes_per_launch[OSQP_CUDA_K_SPMV_A] = bA + 8.0 * m;                 // + rho
  e->bytes_per_launch[OSQP_CUDA_K_SPMV_M] = bM + 8.0 * n;                 // + p (sigma p)
  e->bytes_per_launch[OSQP_CUDA_K_RHS] = bM + 8.0 * 5 * n;                // q, Minv, r, p, xt
  e->bytes_per_launch[OSQP_CUDA_K_PCG_UPDATE] = 8.0 * 7 * n;
  e->bytes_per_launch[OSQP_CUDA_K_PCG_DIR] = 8.0 * 4 * n;
  e->bytes_per_launch[OSQP_CUDA_K_ADMM_UPDATE] = 8.0 * (4 * n) + 8.0 * (14 * m);   // no SpMV: z~ = Ax + zacc (+ w)
  e->bytes_per_launch[OSQP_CUDA_K_RESID_M] = bA + 8.0 * n + 8.0 * 11 * m;
  e->bytes_per_launch[OSQP_CUDA_K_RESID_N] = bM + 8.0 * (n + m) + 8.0 * 5 * n;
  e->bytes_per_launch[OSQP_CUDA_K_TILE_M] = bM;       // the product alone (spmv_tile.cuh, PARTIAL mode: first of two kernels)
  e->bytes_per_launch[OSQP_CUDA_K_TILE_A] = bA;
  e->bytes_per_launch[OSQP_CUDA_K_ALLREDUCE] = 8.0 * n;               // message size (row-partitioned mode)
  e->bytes_per_launch[OSQP_CUDA_K_APPLY_EPI] = 8.0 * 2 * n;
  *ep = e;
  return 0;
}

static int d2h(double *host, const double *dev, size_t len, cudaStream_t st) {
  if (!host || !len) return 0;
  return cudaMemcpyAsync(host, dev, sizeof(double) * len, cudaMemcpyDeviceToHost, st) != cudaSuccess;
}

// K11 -- reference src/scaling.c:44-156 (scale_data)
extern "C" c_int osqp_cuda_engine_scale(osqp_cuda_engine *e, c_int iters, c_float *D, c_float *Dinv,
                                        c_float *E, c_float *Einv, c_float *c, c_float *q_s,
                                        c_float *l_s, c_float *u_s) {
  const int n = e->n, m = e->m;
  e->scaling_iters = (int)iters;
  // restart from the caller's un-scaled data
  if (mats_load_raw(e->W, e->st)) return 1;
  cudaMemcpyAsync(e->q, e->q_raw, sizeof(double) * n, cudaMemcpyDeviceToDevice, e->st);
  engine_reset_scaling(e);
  for (int it = 0; it < (int)iters; ++it) {
    // column inf-norms of [P A'; A 0]: rows of M give D, rows of A give E
    cudaMemsetAsync(e->Dtmp, 0, sizeof(double) * n, e->st);
    mats_row_absmax(e->W.M, e->Dtmp, false, n, e->st);
    if (dist_allreduce(e->dist, e->Dtmp, n, true, false, e->st)) return 1;   // column norms of [P; A] over all slabs
    k_ruiz_vec<<<EW(n)>>>(e->Dtmp, e->D, e->q, n);
    if (m) {
      cudaMemsetAsync(e->Etmp, 0, sizeof(double) * m, e->st);
      mats_row_absmax(e->W.A, e->Etmp, false, n, e->st);
      k_ruiz_vec<<<EW(m)>>>(e->Etmp, e->E, nullptr, m);
    }
    mats_scale(e->W, e->Dtmp, e->Etmp, e->st);
    // cost normalisation
    cudaMemsetAsync(e->Dtmp, 0, sizeof(double) * n, e->st);
    mats_row_absmax(e->W.M, e->Dtmp, true, n, e->st);
    if (dist_allreduce(e->dist, e->Dtmp, n, true, false, e->st)) return 1;   // P's nonzeros are spread over the ranks
    k_cost_scale<<<1, 1024, 0, e->st>>>(e->Dtmp, e->q, n, e->cs);
    mats_scale_P(e->W, e->cs + 2, e->st);
    k_scale_by_dev<<<EW(n)>>>(e->q, e->cs + 2, n);
  }
  mats_sell_sync(e->W, e->st);   // the SpMV mirrors follow the scaled values
  k_recip<<<EW(n)>>>(e->Dinv, e->D, n);
  if (m) {
    k_recip<<<EW(m)>>>(e->Einv, e->E, m);
    k_mul<<<EW(m)>>>(e->l, e->l_raw, e->E, m);
    k_mul<<<EW(m)>>>(e->u, e->u_raw, e->E, m);
  }
  if (getenv("OSQP_B200_TIMING")) { cudaStreamSynchronize(e->st); fprintf(stderr, "[timing]   scale: kernels done, D2H of D, E, q, l, u next\n"); }
  double cs_h[3] = {1.0, 1.0, 1.0};
  int bad = cudaMemcpyAsync(cs_h, e->cs, sizeof(cs_h), cudaMemcpyDeviceToHost, e->st) != cudaSuccess;
  bad |= d2h(D, e->D, n, e->st) | d2h(Dinv, e->Dinv, n, e->st) | d2h(E, e->E, m, e->st) |
         d2h(Einv, e->Einv, m, e->st) | d2h(q_s, e->q, n, e->st) | d2h(l_s, e->l, m, e->st) |
         d2h(u_s, e->u, m, e->st);
  CUDA_TRY(cudaStreamSynchronize(e->st));
  CUDA_TRY(cudaGetLastError());
  if (bad) return 1;
  e->c_host = cs_h[0];
  if (c) *c = cs_h[0];
  e->v_dirty = true;
  return 0;
}

// (re)build rho-dependent state of the linear system; probe = run the curvature test
static c_int engine_refresh_linsys(osqp_cuda_engine *e, bool probe) {
  e->pcg.sigma = e->sigma;
  cudaMemsetAsync(&e->pcg.S->noncvx, 0, sizeof(int), e->st);
  pcg_jacobi(e->pcg, e->L);
  band_refactor(e->band, e->W, e->rho_vec, e->sigma, &e->pcg.S->noncvx, e->st);
  if (e->slack.on) {          // rho~, D and the Schur complement's Jacobi diagonal (slack.cuh)
    SlackDev &sk = e->slack;
    PcgDev &S = e->pcg;
    cudaMemcpyAsync(sk.rho_cg, e->rho_vec, sizeof(double) * e->m, cudaMemcpyDeviceToDevice, e->st);
    k_slack_refresh<<<grid_for(sk.ny), OSQP_CUDA_THREADS, 0, e->st>>>(sk.ny, sk.row, sk.posA, sk.posP, e->W.A.val,
                                                                     e->W.M.val, e->rho_vec, e->sigma, sk.a, sk.pjj,
                                                                     sk.D, sk.rho_cg);
    double *const rho_save = S.rho, *const minv_save = S.Minv;
    S.rho = sk.rho_cg; S.Minv = sk.Minv_cg;
    pcg_jacobi(S, e->L);
    S.rho = rho_save; S.Minv = minv_save;
    k_slack_mask<<<grid_for(sk.ny), OSQP_CUDA_THREADS, 0, e->st>>>(sk.ny, sk.var, sk.Minv_cg);
  }
  e->v_dirty = true;
  if (!probe) return 0;
  // Curvature probe: CG on K z = b for a fixed pseudo-random b; p'Kp <= 0 proves that
  // K = P + sigma I + A' rho A is not positive definite, the same condition the
  // reference reads off the LDL' inertia (qdldl_interface.c:93-98).
  const int n = e->n;
  std::vector<double> b(n);
  unsigned long long s = 0x9E3779B97F4A7C15ull;
  for (int i = 0; i < n; ++i) {
    s = s * 6364136223846793005ull + 1442695040888963407ull;
    b[i] = ((double)(s >> 11) / 9007199254740992.0) - 0.5;
  }
  cudaMemcpyAsync(e->dx, b.data(), sizeof(double) * n, cudaMemcpyHostToDevice, e->st);
  int conv = 0;
  const int iters = pcg_solve_sync(e->pcg, e->L, e->dx, e->xt, true, 1e-12, n < 30 ? n + 2 : 30, &conv);
  cudaMemsetAsync(e->dx, 0, sizeof(double) * n, e->st);
  PcgScal h;
  CUDA_TRY(cudaMemcpyAsync(&h, e->pcg.S, sizeof(PcgScal), cudaMemcpyDeviceToHost, e->st));
  CUDA_TRY(cudaStreamSynchronize(e->st));
  if (iters < 0) return OSQP_LINSYS_SOLVER_INIT_ERROR;
  if (h.noncvx) return OSQP_NONCVX_ERROR;
  return 0;
}

static void plugin_bind_engine(osqp_cuda_engine *e);   // defined with the plugin below

// K5 -- reference src/auxil.c:76-98 (set_rho_vec) + linear-system initialisation
extern "C" c_int osqp_cuda_engine_init_linsys(osqp_cuda_engine *e, c_float sigma, c_float rho,
                                              cuda_pcg_solver **solver_out) {
  e->sigma = sigma; e->rho = rho;
  if (e->m) k_rho_types<<<EW(e->m)>>>(e->l, e->u, rho, e->m, e->ctype, e->rho_vec, 0, e->flag);
  const c_int rc = engine_refresh_linsys(e, true);
  if (rc) return rc;
  e->linsys_ready = true;
  plugin_bind_engine(e);
  if (solver_out) *solver_out = e->solver;
  return 0;
}

// reference src/osqp.c:1302-1323
extern "C" c_int osqp_cuda_engine_update_rho(osqp_cuda_engine *e, c_float rho) {
  e->rho = rho;
  if (e->m) k_rho_update<<<EW(e->m)>>>(e->ctype, rho, e->m, e->rho_vec);
  e->stats.rho_updates += 1;
  if (e->pcg_probe < 3) e->pcg_probe = 3;
  return engine_refresh_linsys(e, false);
}

// reference src/auxil.c:100-142 (update_rho_vec)
extern "C" c_int osqp_cuda_engine_retype_constraints(osqp_cuda_engine *e, c_float rho, c_int *changed) {
  int h = 0;
  e->rho = rho;
  if (e->m) {
    cudaMemsetAsync(e->flag, 0, sizeof(int), e->st);
    k_rho_types<<<EW(e->m)>>>(e->l, e->u, rho, e->m, e->ctype, e->rho_vec, 1, e->flag);
    if (dist_allreduce(e->dist, e->flag, 1, true, true, e->st)) return 1;   // any slab changed -> all refresh
    CUDA_TRY(cudaMemcpyAsync(&h, e->flag, sizeof(int), cudaMemcpyDeviceToHost, e->st));
    CUDA_TRY(cudaStreamSynchronize(e->st));
  }
  if (changed) *changed = h;
  if (h) return engine_refresh_linsys(e, false);
  return 0;
}

extern "C" c_int osqp_cuda_engine_set_pcg_lambda(osqp_cuda_engine *e, c_float lambda) {
  e->pcg_lambda = lambda;      // <= 0 disables the residual-based cap
  return 0;
}
extern "C" c_int osqp_cuda_engine_set_pcg(osqp_cuda_engine *e, c_float rel_tol, c_float abs_floor,
                                          c_int max_iter) {
  if (rel_tol > 0) e->pcg_rel = rel_tol;
  if (abs_floor > 0) e->pcg_abs = abs_floor;
  if (max_iter > 0) e->pcg_max = (int)max_iter;
  return 0;
}

// Ax <- A x, v <- y + rho .* (Ax - z) after anything but the ADMM loop touched x, z, y or rho
static void engine_refresh_v(osqp_cuda_engine *e) {
  if (e->m) {
    EpiAxV ep{e->rho_vec, e->z, e->y, e->Ax, e->v};
    launch_stream<1, false, 0>(e->L, -1, e->W.A, e->pcg.gridA, Gather{e->x, nullptr},
                               Gather{nullptr, nullptr}, ep, nullptr);
  }
  e->v_dirty = false;
}

static void engine_enqueue_one(osqp_cuda_engine *e, double alpha, int store, int pcg_count) {
  PcgDev &S = e->pcg;
  // slack variables eliminated from the CG's system (slack.cuh): the CG runs on the Schur complement
  // (rho~, inverse diagonal 0 on the slacks); their part of the rhs is folded into v first
  SlackDev &sk = e->slack;
  double *const rho_save = S.rho, *const minv_save = S.Minv;
  if (sk.on) {
    e->L.pre(-1);
    k_slack_pre<<<grid_for(sk.ny), OSQP_CUDA_THREADS, 0, e->st>>>(sk.ny, sk.row, sk.var, sk.a, sk.pjj, sk.D,
                                                                 e->rho_vec, e->q, e->x, e->v, sk.r0);
    e->L.post();
    S.rho = sk.rho_cg; S.Minv = sk.Minv_cg;
  }
  // K_a: r0, p0, x~ <- x
  EpiRhs er{S.S, e->q, e->x, S.Minv, S.r, S.p, e->xt};
  launch_stream<1, false, 0>(e->L, OSQP_CUDA_K_RHS, e->W.M, S.gridM, Gather{e->x, e->v},
                             Gather{nullptr, nullptr}, er, S.partRZ[0]);
  // K4: CG on x~ (accumulating A (x~ - x) on the side)
  const int partsM = stream_parts(e->L, e->W.M, S.gridM);
  pcg_enqueue(S, e->L, e->xt, 0, pcg_count, partsM, e->zacc);
  // K_f: close the CG, x update, z~ = Ax + A (x~ - x), z / y / Ax / v updates
  {
    const int len = e->n > e->m ? e->n : e->m;
    e->L.pre(OSQP_CUDA_K_ADMM_UPDATE);
    k_admm_update<<<EW(len)>>>(S.S, S.partRZ[pcg_count & 1], pcg_count == 0 ? partsM : S.grid_n, e->n, e->m,
                               alpha, store, e->x, e->xt, e->dx, e->rho_vec, e->l, e->u, e->z, e->y, e->Ax,
                               e->v, e->dy, S.w, e->zacc, slack_view(sk));
    e->L.post();
  }
  S.rho = rho_save; S.Minv = minv_save;
  e->stats.admm_iters += 1;
}

static void engine_drop_graphs(osqp_cuda_engine *e) {
  for (auto &g : e->graphs) cudaGraphExecDestroy(g.exec);
  e->graphs.clear();
}

// Replays (capturing on first use) the graph of one ADMM iteration; false = caller enqueues directly.
static bool engine_graph_one(osqp_cuda_engine *e, double alpha, int store, int pcg_count) {
  PcgDev &S = e->pcg;
  if (!e->graphs_on || e->L.profiling || !e->warmed) return false;
  for (auto &g : e->graphs) {
    if (g.L == pcg_count && g.store == store && g.alpha == alpha && g.rel == S.rel_tol &&
        g.absf == S.abs_floor && g.maxit == S.max_iter && g.use_cap == S.use_cap &&
        g.gridA == S.gridA && g.gridM == S.gridM) {
      if (cudaGraphLaunch(g.exec, e->st) != cudaSuccess) return false;
      e->L.launches += g.launches;
      e->stats.admm_iters += 1;
      return true;
    }
  }
  if (e->graphs.size() >= 24) engine_drop_graphs(e);
  osqp_cuda_engine::IterGraph g{pcg_count, store, S.max_iter, S.use_cap, S.gridA, S.gridM,
                                alpha, S.rel_tol, S.abs_floor, 0, nullptr};
  const long long l0 = e->L.launches;
  const c_int it0 = e->stats.admm_iters;
  if (cudaStreamBeginCapture(e->st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
    cudaGetLastError();
    e->graphs_on = false;
    return false;
  }
  engine_enqueue_one(e, alpha, store, pcg_count);
  cudaGraph_t graph = nullptr;
  const cudaError_t ce = cudaStreamEndCapture(e->st, &graph);
  g.launches = e->L.launches - l0;
  e->L.launches = l0;
  e->stats.admm_iters = it0;
  if (ce != cudaSuccess || !graph || cudaGraphInstantiate(&g.exec, graph, 0) != cudaSuccess) {
    if (graph) cudaGraphDestroy(graph);
    cudaGetLastError();
    e->graphs_on = false;      // fall back to plain launches for the rest of this engine's life
    return false;
  }
  cudaGraphDestroy(graph);
  e->graphs.push_back(g);
  if (cudaGraphLaunch(g.exec, e->st) != cudaSuccess) return false;
  e->L.launches += g.launches;
  e->stats.admm_iters += 1;
  return true;
}


// Spare CG iterations enqueued over the largest count of the last interval (OSQP_B200_PCG_SPARE).  Default
// 0: a spare is 5 no-op launches per ADMM iteration (measured: config 2 3956 -> 4040 it/s without it, same
// iteration count); a CG that needs one more than the last interval's maximum is cut short -- a valid
// inexact ADMM step -- and the policy grows L and re-probes.
static int pcg_spare() {
  static int v = -1;
  if (v < 0) { const char *e = getenv("OSQP_B200_PCG_SPARE"); v = e ? atoi(e) : 0; if (v < 0) v = 0; }
  return v;
}
// CG settings of the ADMM solves of this engine (read by the iteration kernels)
static void engine_prepare_pcg(osqp_cuda_engine *e) {
  PcgDev &S = e->pcg;
  S.rel_tol = e->pcg_rel; S.abs_floor = e->pcg_abs; S.max_iter = e->pcg_max; S.use_cap = 1;
  if (e->n <= OSQP_PCG_EXACT_N) {
    // tiny systems: CG terminates in <= n steps, so the linear solves are run to (near)
    // machine precision and the ADMM trajectory is that of the reference's direct solver
    S.rel_tol = e->pcg_rel < 1e-12 ? e->pcg_rel : 1e-12;
    S.use_cap = 0;
    S.max_iter = 2 * e->n + 8;
    e->pcg_launch = S.max_iter;
    e->pcg_probe = 0;
  }
}
// Small problems: the single-CTA kernels walk the sliced-ELL mirrors when both exist and are current
static int engine_use_sell_rows(const osqp_cuda_engine *e) {
  const SellDev &a = e->W.A.sell, &m = e->W.M.sell;
  return m.ok && !m.dirty && (e->m == 0 || e->W.A.nnz == 0 || (a.ok && !a.dirty));
}
// Arguments of the fused single-CTA kernel for `count` iterations of this engine
static FusedArgs engine_fused_args(osqp_cuda_engine *e, double alpha, int count, int store_delta) {
  PcgDev &S = e->pcg;
  FusedArgs a;
  a.A = view_of(e->W.A); a.M = view_of(e->W.M);
  a.As = sell_view_of(e->W.A); a.Ms = sell_view_of(e->W.M);
  a.use_sell = engine_use_sell_rows(e);
  a.n = e->n; a.m = e->m;
  a.x = e->x; a.z = e->z; a.y = e->y; a.Ax = e->Ax; a.v = e->v; a.dx = e->dx; a.dy = e->dy;
  a.q = e->q; a.l = e->l; a.u = e->u; a.rho = e->rho_vec; a.Minv = S.Minv;
  a.S = S.S;
  a.sigma = S.sigma; a.alpha = alpha; a.rel = S.rel_tol; a.absf = S.abs_floor;
  a.maxit = S.max_iter; a.use_cap = S.use_cap; a.count = count; a.store = store_delta;
  a.use_band = e->band.ok ? 1 : 0;
  a.band = band_view_of(e->band);
  a.smem_bytes = (unsigned)(admm_fused_smem(e->n, e->m) + band_apply_smem(e->band));
  return a;
}

// reference src/osqp.c:354-370 -- `count` ADMM iterations, enqueued without host sync
extern "C" c_int osqp_cuda_engine_iterate(osqp_cuda_engine *e, c_int count, c_float alpha,
                                          c_int store_delta) {
  if (!e->linsys_ready) return 1;
  engine_prepare_pcg(e);
  PcgDev &S = e->pcg;
  if (e->v_dirty) engine_refresh_v(e);
  if (e->fused) {
    // one launch = `count` complete ADMM iterations (admm_fused.cuh); the CG loop runs to its
    // stopping rule on the device, so there is no launch policy and nothing to probe
    FusedArgs a = engine_fused_args(e, alpha, (int)count, (int)store_delta);
    e->stats.admm_iters += count;
    e->L.pre(OSQP_CUDA_K_ADMM_UPDATE);
    admm_fused_kernel<<<1, ADMM_FUSED_THREADS, a.smem_bytes, e->st>>>(a);
    e->L.post();
    return cudaGetLastError() != cudaSuccess;
  }
  c_int k = 0;
  // Probe mode (start of a solve, after a rho update or after a truncated CG): ADMM
  // iterations are enqueued one at a time and the CG iteration count is read back, so
  // the number of CG iterations launched per ADMM iteration tracks what the stopping
  // rule needs.  A truncated CG is still a valid (inexact) ADMM step; it is not re-run.
  while (e->pcg_probe > 0 && k < count) {
    engine_enqueue_one(e, alpha, (k == count - 1) ? (int)store_delta : 0, e->pcg_launch);
    e->warmed = true;
    PcgScal h;
    CUDA_TRY(cudaMemcpyAsync(&h, S.S, sizeof(PcgScal), cudaMemcpyDeviceToHost, e->st));
    CUDA_TRY(cudaStreamSynchronize(e->st));
    if (!h.conv && h.iters >= e->pcg_launch) {
      e->pcg_launch = 2 * e->pcg_launch + 2;
      if (e->pcg_probe < 2) e->pcg_probe = 2;
    } else {
      // `want` enqueued iterations can apply `want` updates (the closing kernel runs the last
      // convergence test): what was needed (+ pcg_spare()), a quarter more for long solves.
      const int want = h.iters + pcg_spare() + h.iters / 4;
      // shrink slowly, grow at once
      e->pcg_launch = want > e->pcg_launch ? want : (want + 3 * e->pcg_launch) / 4;
      e->pcg_probe -= 1;
    }
    if (e->pcg_launch < 2) e->pcg_launch = 2;
    if (e->pcg_launch > S.max_iter) e->pcg_launch = S.max_iter;
    ++k;
  }
  for (; k < count; ++k) {
    const int store = (k == count - 1) ? (int)store_delta : 0;
    if (!engine_graph_one(e, alpha, store, e->pcg_launch)) {
      engine_enqueue_one(e, alpha, store, e->pcg_launch);
      e->warmed = true;
    }
  }
  return cudaGetLastError() != cudaSuccess;
}

// Host side of a residual pass: the scalars `h` (layout of k_res_final) -> osqp_cuda_resid, and
// the CG statistics of the interval -> launch policy of the next one.
static void resid_fill(osqp_cuda_engine *e, const double *h, osqp_cuda_resid *out, int polished) {
  const int m = e->m;
  if (!polished) e->v_dirty = false;
  const double *hm = h + RESM_NSUM, *hn = h + (RESM_NSUM + RESM_NMAX) + RESN_NSUM;
  const double *sn = h + (RESM_NSUM + RESM_NMAX);
  memset(out, 0, sizeof(*out));
  if (m) {
    out->ineq_lhs = h[0];
    out->pri_res_s = hm[0]; out->pri_res_u = hm[1];
    out->z_norm_s = hm[2]; out->z_norm_u = hm[3];
    out->Ax_norm_s = hm[4]; out->Ax_norm_u = hm[5];
    out->dy_norm_s = hm[6] < 0 ? 0 : hm[6]; out->dy_norm_u = hm[7] < 0 ? 0 : hm[7];
    out->Adx_maxU_s = hm[8]; out->Adx_maxU_u = hm[9];
    out->Adx_minL_s = -hm[10]; out->Adx_minL_u = -hm[11];
  } else {
    out->Adx_maxU_s = out->Adx_maxU_u = -DBL_MAX;
    out->Adx_minL_s = out->Adx_minL_u = DBL_MAX;
  }
  out->xPx = sn[0]; out->qx = sn[1]; out->q_dx = sn[2];
  out->dua_res_s = hn[0]; out->dua_res_u = hn[1];
  out->q_norm_s = hn[2]; out->q_norm_u = hn[3];
  out->Aty_norm_s = hn[4]; out->Aty_norm_u = hn[5];
  out->Px_norm_s = hn[6]; out->Px_norm_u = hn[7];
  out->Atdy_norm_s = hn[8]; out->Atdy_norm_u = hn[9];
  out->dx_norm_s = hn[10]; out->dx_norm_u = hn[11];
  out->Pdx_norm_s = hn[12]; out->Pdx_norm_u = hn[13];
  // PCG statistics of the interval -> launch policy for the next one
  const double *ps = h + (RESM_NSUM + RESM_NMAX + RESN_NSUM + RESN_NMAX);
  e->stats.pcg_iters = (c_int)ps[0];
  const int maxed = (int)ps[2], max_it = (int)ps[3];
  e->stats.pcg_maxed += maxed;
  e->stats.pcg_breakdowns = (c_int)ps[5];
  e->stats.pcg_noncvx = ps[4] != 0.0;
  out->pcg_noncvx = ps[4];
  if (!polished && e->n > OSQP_PCG_EXACT_N) {
    if (maxed > 0) {
      e->pcg_launch = 2 * e->pcg_launch + 2;
      if (e->pcg_probe < 2) e->pcg_probe = 2;
    } else if (max_it > 0) {
      const int want = max_it + pcg_spare() + max_it / 4;
      e->pcg_launch = want > e->pcg_launch ? want : (want + e->pcg_launch) / 2;
    }
    if (e->pcg_launch > e->pcg_max) e->pcg_launch = e->pcg_max;
    if (e->pcg_launch < 2) e->pcg_launch = 2;
  }
}

// K9 + K10 -- reference src/auxil.c:240-359 (residuals, tolerances), :361-512 (infeasibility)
extern "C" c_int osqp_cuda_engine_residuals(osqp_cuda_engine *e, osqp_cuda_resid *out, c_int polished) {
  const int n = e->n, m = e->m;
  PcgDev &S = e->pcg;
  const double *x = polished ? e->pol_x : e->x;
  const double *z = polished ? e->pol_z : e->z;
  const double *y = polished ? e->pol_y : e->y;
  int gM = 0;
  if (m) {
    EpiResM em{z, y, e->rho_vec, e->l, e->u, e->E, e->Einv, e->dy, e->Ax, e->v, e->dyp,
               polished ? 0 : 1};
    gM = S.gridA;
    launch_stream<2, false, 0>(e->L, OSQP_CUDA_K_RESID_M, e->W.A, gM, Gather{x, nullptr},
                               Gather{e->dx, nullptr}, em, e->partM);
  }
  EpiResN en{e->q, x, e->dx, e->D, e->Dinv};
  launch_stream<2, true, 0>(e->L, OSQP_CUDA_K_RESID_N, e->W.M, S.gridM, Gather{x, y},
                            Gather{e->dx, e->dyp}, en, e->partN);
  const double *partM = e->partM;
  if (e->dist && m) {
    // m-side scalars of the slabs: compact the per-CTA partials, join sums and maxes
    k_compact_partials<<<1, OSQP_CUDA_THREADS, 0, e->st>>>(e->partM, gM, RESM_NSUM, RESM_NMAX, e->partMc);
    if (dist_allreduce(e->dist, e->partMc, RESM_NSUM, false, false, e->st) ||
        dist_allreduce(e->dist, e->partMc + RESM_NSUM, RESM_NMAX, true, false, e->st)) e->L.failed = true;
    partM = e->partMc;
    gM = 1;
  }
  e->L.pre(-1);
  k_res_final<<<1, OSQP_CUDA_THREADS, 0, e->st>>>(partM, gM, e->partN, S.gridM, S.S, e->res_dev,
                                                  e->pcg_lambda, polished ? 0 : 1);
  e->L.post();
  CUDA_TRY(cudaMemcpyAsync(e->res_host, e->res_dev, sizeof(double) * RES_OUT_LEN,
                           cudaMemcpyDeviceToHost, e->st));
  CUDA_TRY(cudaStreamSynchronize(e->st));
  CUDA_TRY(cudaGetLastError());
  if (e->L.failed) {
    fprintf(stderr, "osqp_cuda_pcg: a collective of the row-partitioned mode failed; the iterates are invalid\n");
    return 1;
  }
  resid_fill(e, e->res_host, out, (int)polished);
  return 0;
}

// ------------------------------------------------ lockstep batches of small problems
// osqp_solve_batch on fused-size problems (BASELINE config 4): ONE host thread drives all
// problems through the same check interval with two launches for the whole batch -- the fused
// ADMM kernel with a CTA per problem (admm_fused_batch_kernel) and the residual kernel below,
// also a CTA per problem -- one D2H copy of all residual scalars and one synchronisation.  The
// per-problem host logic (check_termination, adapt_rho: osqp_api.c) is unchanged; what goes
// away is ~10 driver calls per problem and interval, which is what bounded the threaded batch.
struct BatchResArgs {
  CsrView A, M;
  SellView As, Ms;
  int use_sell;
  int m;
  EpiResM em;
  EpiResN en;
  Gather gA0, gA1, gM0, gM1;
  PcgScal *S;
  double *out;            // RES_OUT_LEN doubles, layout of k_res_final
  double lambda;
};
// Same row epilogues as osqp_cuda_engine_residuals; the per-CTA partials + k_res_final become
// CTA-wide reductions (max-norms are order independent; the three sums differ in the last bit).
__global__ void __launch_bounds__(OSQP_CUDA_THREADS) resid_batch_kernel(const BatchResArgs *args) {
  __shared__ BatchResArgs a;
  __shared__ double scratch[32];
  if (threadIdx.x == 0) a = args[blockIdx.x];
  __syncthreads();
  const int warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int KM = RESM_NSUM + RESM_NMAX, KN = RESN_NSUM + RESN_NMAX;
  double *out = a.out;
  {
    double asum[RESM_NSUM], amax[RESM_NMAX];
#pragma unroll
    for (int k = 0; k < RESM_NSUM; ++k) asum[k] = 0.0;
#pragma unroll
    for (int k = 0; k < RESM_NMAX; ++k) amax[k] = -DBL_MAX;
    if (a.m) {
      EpiResM em = a.em;
      if (a.use_sell) spmv_sell_rows<2, false, 0>(a.As, a.gA0, a.gA1, em, asum, amax, warp, nwarps);
      else spmv_direct_rows<2, false, 0, false>(a.A, a.gA0, a.gA1, em, asum, amax, warp, nwarps);
    }
#pragma unroll
    for (int k = 0; k < RESM_NSUM; ++k) {
      const double t = block_reduce<false>(asum[k], scratch);
      if (threadIdx.x == 0) out[k] = t;
    }
#pragma unroll
    for (int k = 0; k < RESM_NMAX; ++k) {
      const double t = block_reduce<true>(amax[k], scratch);
      if (threadIdx.x == 0) out[RESM_NSUM + k] = t;
    }
  }
  __syncthreads();      // the projected delta_y written by the m-side rows is gathered below
  {
    double asum[RESN_NSUM], amax[RESN_NMAX];
#pragma unroll
    for (int k = 0; k < RESN_NSUM; ++k) asum[k] = 0.0;
#pragma unroll
    for (int k = 0; k < RESN_NMAX; ++k) amax[k] = -DBL_MAX;
    EpiResN en = a.en;
    if (a.use_sell) spmv_sell_rows<2, true, 0>(a.Ms, a.gM0, a.gM1, en, asum, amax, warp, nwarps);
    else spmv_direct_rows<2, true, 0, false>(a.M, a.gM0, a.gM1, en, asum, amax, warp, nwarps);
#pragma unroll
    for (int k = 0; k < RESN_NSUM; ++k) {
      const double t = block_reduce<false>(asum[k], scratch);
      if (threadIdx.x == 0) out[KM + k] = t;
    }
#pragma unroll
    for (int k = 0; k < RESN_NMAX; ++k) {
      const double t = block_reduce<true>(amax[k], scratch);
      if (threadIdx.x == 0) out[KM + RESN_NSUM + k] = t;
    }
  }
  if (threadIdx.x == 0) {
    PcgScal *S = a.S;
    out[KM + KN + 0] = (double)S->total_iters;
    out[KM + KN + 1] = (double)S->solves;
    out[KM + KN + 2] = (double)S->maxed;
    out[KM + KN + 3] = (double)S->max_iters;
    out[KM + KN + 4] = (double)S->noncvx;
    out[KM + KN + 5] = (double)S->breakdowns;
    S->maxed = 0;
    S->max_iters = 0;
    const double pri = a.m ? out[RESM_NSUM + 0] : 0.0, dua = out[KM + RESN_NSUM + 0];
    double cap = (pri > 0.0) ? a.lambda * sqrt(pri * dua) : a.lambda * dua;
    if (!(a.lambda > 0.0) || !(cap == cap)) cap = DBL_MAX;
    S->tol_cap = cap;
  }
}

struct BatchStep {        // staging buffers of osqp_cuda_batch_step (one caller at a time)
  pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
  cudaStream_t st = nullptr;
  int cap = 0;
  FusedArgs *hf = nullptr, *df = nullptr;
  BatchResArgs *hr = nullptr, *dr = nullptr;
  double *hout = nullptr, *dout = nullptr;
};
static BatchStep g_bs;

extern "C" c_int osqp_cuda_engine_is_fused(const osqp_cuda_engine *e) { return e && e->fused && e->linsys_ready; }

// `count` ADMM iterations on every engine of the batch, then (do_resid) the residual pass of
// each, results in outs[i].  All engines must be fused-size (osqp_cuda_engine_is_fused).
extern "C" c_int osqp_cuda_batch_step(osqp_cuda_engine **es, c_int n, c_int count, c_float alpha,
                                      c_int store_delta, c_int do_resid, osqp_cuda_resid **outs) {
  if (n <= 0) return 0;
  if (bind_device()) return 1;
  BatchStep &B = g_bs;
  pthread_mutex_lock(&B.mu);
  int rc = 1;
  do {
    if (!B.st && cudaStreamCreateWithFlags(&B.st, cudaStreamNonBlocking) != cudaSuccess) break;
    if (n > B.cap) {
      if (B.hf) { cudaFreeHost(B.hf); cudaFreeHost(B.hr); cudaFreeHost(B.hout); cudaFree(B.df); cudaFree(B.dr); cudaFree(B.dout); }
      B.hf = nullptr; B.cap = 0;
      const size_t c = (size_t)n + 64;
      if (cudaMallocHost((void **)&B.hf, c * sizeof(FusedArgs)) != cudaSuccess ||
          cudaMallocHost((void **)&B.hr, c * sizeof(BatchResArgs)) != cudaSuccess ||
          cudaMallocHost((void **)&B.hout, c * RES_OUT_LEN * sizeof(double)) != cudaSuccess ||
          cudaMalloc((void **)&B.df, c * sizeof(FusedArgs)) != cudaSuccess ||
          cudaMalloc((void **)&B.dr, c * sizeof(BatchResArgs)) != cudaSuccess ||
          cudaMalloc((void **)&B.dout, c * RES_OUT_LEN * sizeof(double)) != cudaSuccess) break;
      B.cap = (int)c;
    }
    size_t smem = 0;
    bool ok = true;
    for (int i = 0; i < n && ok; ++i) {
      osqp_cuda_engine *e = es[i];
      if (!e || !e->fused || !e->linsys_ready) { ok = false; break; }
      engine_prepare_pcg(e);
      if (e->v_dirty) engine_refresh_v(e);
      B.hf[i] = engine_fused_args(e, alpha, (int)count, (int)store_delta);
      smem = std::max(smem, (size_t)B.hf[i].smem_bytes);
      if (do_resid) {
        BatchResArgs &r = B.hr[i];
        r.A = view_of(e->W.A); r.M = view_of(e->W.M); r.m = e->m;
        r.As = sell_view_of(e->W.A); r.Ms = sell_view_of(e->W.M); r.use_sell = engine_use_sell_rows(e);
        r.em = EpiResM{e->z, e->y, e->rho_vec, e->l, e->u, e->E, e->Einv, e->dy, e->Ax, e->v, e->dyp, 1};
        r.en = EpiResN{e->q, e->x, e->dx, e->D, e->Dinv};
        r.gA0 = Gather{e->x, nullptr}; r.gA1 = Gather{e->dx, nullptr};
        r.gM0 = Gather{e->x, e->y};    r.gM1 = Gather{e->dx, e->dyp};
        r.S = e->pcg.S; r.out = B.dout + (size_t)i * RES_OUT_LEN; r.lambda = e->pcg_lambda;
      }
    }
    if (!ok) break;
    // per-engine work enqueued so far (cold start, v refresh, rho updates) precedes the batch
    if (cudaDeviceSynchronize() != cudaSuccess) break;
    cudaMemcpyAsync(B.df, B.hf, (size_t)n * sizeof(FusedArgs), cudaMemcpyHostToDevice, B.st);
    admm_fused_batch_kernel<<<(int)n, ADMM_FUSED_THREADS, smem, B.st>>>(B.df);
    if (do_resid) {
      cudaMemcpyAsync(B.dr, B.hr, (size_t)n * sizeof(BatchResArgs), cudaMemcpyHostToDevice, B.st);
      resid_batch_kernel<<<(int)n, OSQP_CUDA_THREADS, 0, B.st>>>(B.dr);
      cudaMemcpyAsync(B.hout, B.dout, (size_t)n * RES_OUT_LEN * sizeof(double), cudaMemcpyDeviceToHost, B.st);
    }
    if (cudaStreamSynchronize(B.st) != cudaSuccess || cudaGetLastError() != cudaSuccess) {
      fprintf(stderr, "osqp_cuda_pcg: batch step failed: %s\n", cudaGetErrorString(cudaGetLastError()));
      break;
    }
    for (int i = 0; i < n; ++i) {
      osqp_cuda_engine *e = es[i];
      e->stats.admm_iters += count;
      e->L.launches += do_resid ? 2 : 1;     // this problem's CTA of each batch kernel
      if (do_resid) resid_fill(e, B.hout + (size_t)i * RES_OUT_LEN, outs[i], 0);
    }
    rc = 0;
  } while (0);
  pthread_mutex_unlock(&B.mu);
  return rc;
}

// reference src/auxil.c:155-159
extern "C" c_int osqp_cuda_engine_cold_start(osqp_cuda_engine *e) {
  cudaSetDevice(g_device);     // may be the first call of an osqp_solve_batch worker thread
  const size_t N = e->n, M = e->m ? e->m : 1;
  cudaMemsetAsync(e->x, 0, sizeof(double) * N, e->st);
  double *zm[] = {e->z, e->y, e->Ax, e->v};
  for (double *p : zm) cudaMemsetAsync(p, 0, sizeof(double) * M, e->st);
  e->v_dirty = false;     // Ax = 0 and v = y + rho (Ax - z) = 0 hold exactly
  e->pcg_probe = 6;
  return 0;
}
// Called by osqp_solve at the start of every solve: same launch policy each time.
extern "C" c_int osqp_cuda_engine_begin_solve(osqp_cuda_engine *e) {
  cudaSetDevice(g_device);     // the calling thread may be a fresh osqp_solve_batch worker
  e->pcg_probe = 6;
  e->pcg_launch = 6;
  pcg_set_cap_kernel<<<1, 1, 0, e->st>>>(e->pcg.S, DBL_MAX);   // no residuals known yet
  // interval statistics left by setup's curvature probe or a polish must not steer the launch policy
  pcg_reset_interval_kernel<<<1, 1, 0, e->st>>>(e->pcg.S);
  return 0;
}

// osqp_cuda_engine_cold_start (if cold) + osqp_cuda_engine_begin_solve for every engine of a batch in
// ONE launch (a CTA per problem) instead of 7 enqueues per problem: 168 ms of a 524 ms batch of 4096
// MPC problems was spent there.  Same stores as the per-engine calls, so the solves stay bitwise equal.
struct BatchBeginArgs { double *x, *z, *y, *Ax, *v; PcgScal *S; int n, m, cold, pad; };
__global__ void __launch_bounds__(OSQP_CUDA_THREADS) batch_begin_kernel(const BatchBeginArgs *args) {
  const BatchBeginArgs a = args[blockIdx.x];
  if (a.cold) {
    for (int i = threadIdx.x; i < a.n; i += blockDim.x) a.x[i] = 0.0;
    const int M = a.m ? a.m : 1;
    for (int i = threadIdx.x; i < M; i += blockDim.x) { a.z[i] = 0.0; a.y[i] = 0.0; a.Ax[i] = 0.0; a.v[i] = 0.0; }
  }
  if (threadIdx.x == 0) { a.S->tol_cap = DBL_MAX; a.S->maxed = 0; a.S->max_iters = 0; }
}
struct BatchUnscaleArgs { const double *x, *D, *y, *E, *cs; int n, m; long long off; };
__global__ void __launch_bounds__(OSQP_CUDA_THREADS) batch_unscale_kernel(const BatchUnscaleArgs *args, double *stage) {
  const BatchUnscaleArgs a = args[blockIdx.x];
  double *out = stage + a.off;
  for (int i = threadIdx.x; i < a.n; i += blockDim.x) out[i] = a.x[i] * a.D[i];          // k_mul
  if (a.m) {
    const double s = a.cs[1];
    for (int i = threadIdx.x; i < a.m; i += blockDim.x) out[a.n + i] = (a.y[i] * a.E[i]) * s;   // k_mul_scal
  }
}
// pinned + device staging for the two kernels above (guarded by g_bs.mu)
static void *g_bb_host = nullptr, *g_bb_dev = nullptr;
static size_t g_bb_cap = 0;
static int batch_args_reserve(size_t bytes) {
  if (bytes <= g_bb_cap) return 0;
  if (g_bb_host) { cudaFreeHost(g_bb_host); cudaFree(g_bb_dev); g_bb_host = g_bb_dev = nullptr; g_bb_cap = 0; }
  bytes += 4096;
  if (cudaMallocHost(&g_bb_host, bytes) != cudaSuccess || cudaMalloc(&g_bb_dev, bytes) != cudaSuccess) return 1;
  g_bb_cap = bytes;
  return 0;
}
extern "C" c_int osqp_cuda_batch_begin_solve(osqp_cuda_engine **es, c_int n, c_int cold) {
  if (n <= 0) return 0;
  if (bind_device()) return 1;
  BatchStep &B = g_bs;
  pthread_mutex_lock(&B.mu);
  int rc = 1;
  do {
    if (!B.st && cudaStreamCreateWithFlags(&B.st, cudaStreamNonBlocking) != cudaSuccess) break;
    if (batch_args_reserve((size_t)n * sizeof(BatchBeginArgs))) break;
    BatchBeginArgs *h = (BatchBeginArgs *)g_bb_host;
    for (c_int i = 0; i < n; ++i) {
      osqp_cuda_engine *e = es[i];
      h[i] = BatchBeginArgs{e->x, e->z, e->y, e->Ax, e->v, e->pcg.S, e->n, e->m, (int)cold, 0};
      if (cold) e->v_dirty = false;
      e->pcg_probe = 6; e->pcg_launch = 6;
    }
    // earlier per-engine work (setup, updates) is on the engines' own streams
    if (cudaDeviceSynchronize() != cudaSuccess) break;
    cudaMemcpyAsync(g_bb_dev, h, (size_t)n * sizeof(BatchBeginArgs), cudaMemcpyHostToDevice, B.st);
    batch_begin_kernel<<<(int)n, OSQP_CUDA_THREADS, 0, B.st>>>((const BatchBeginArgs *)g_bb_dev);
    if (cudaStreamSynchronize(B.st) != cudaSuccess || cudaGetLastError() != cudaSuccess) break;
    rc = 0;
  } while (0);
  pthread_mutex_unlock(&B.mu);
  return rc;
}

// reference src/osqp.c:942-1007
extern "C" c_int osqp_cuda_engine_warm_start(osqp_cuda_engine *e, const c_float *x, const c_float *y) {
  const int n = e->n, m = e->m;
  if (x) {
    CUDA_TRY(cudaMemcpyAsync(e->xt, x, sizeof(double) * n, cudaMemcpyHostToDevice, e->st));
    k_mul<<<EW(n)>>>(e->x, e->xt, e->Dinv, n);
    if (m) {
      EpiStore es{e->z};
      launch_stream<1, false, 0>(e->L, -1, e->W.A, e->pcg.gridA, Gather{e->x, nullptr},
                                 Gather{nullptr, nullptr}, es, nullptr);
    }
  }
  if (y && m) {
    CUDA_TRY(cudaMemcpyAsync(e->dy, y, sizeof(double) * m, cudaMemcpyHostToDevice, e->st));
    k_mul_scal<<<EW(m)>>>(e->y, e->dy, e->Einv, e->cs, m);
  }
  e->v_dirty = true;
  CUDA_TRY(cudaStreamSynchronize(e->st));
  return 0;
}

// reference src/osqp.c:765-795
extern "C" c_int osqp_cuda_engine_update_q(osqp_cuda_engine *e, const c_float *q_new, c_float *q_s) {
  const int n = e->n;
  CUDA_TRY(cudaMemcpyAsync(e->q_raw, q_new, sizeof(double) * n, cudaMemcpyHostToDevice, e->st));
  k_mul_scal<<<EW(n)>>>(e->q, e->q_raw, e->D, e->cs, n);
  d2h(q_s, e->q, n, e->st);
  CUDA_TRY(cudaStreamSynchronize(e->st));
  return 0;
}
// reference src/osqp.c:797-940 (the l <= u validation stays on the host)
extern "C" c_int osqp_cuda_engine_update_bounds(osqp_cuda_engine *e, const c_float *l_new,
                                                const c_float *u_new, c_float *l_s, c_float *u_s) {
  const int m = e->m;
  if (!m) return 0;
  if (l_new) {
    CUDA_TRY(cudaMemcpyAsync(e->l_raw, l_new, sizeof(double) * m, cudaMemcpyHostToDevice, e->st));
    k_mul<<<EW(m)>>>(e->l, e->l_raw, e->E, m);
    d2h(l_s, e->l, m, e->st);
  }
  if (u_new) {
    CUDA_TRY(cudaMemcpyAsync(e->u_raw, u_new, sizeof(double) * m, cudaMemcpyHostToDevice, e->st));
    k_mul<<<EW(m)>>>(e->u, e->u_raw, e->E, m);
    d2h(u_s, e->u, m, e->st);
  }
  CUDA_TRY(cudaStreamSynchronize(e->st));
  return 0;
}

// reference src/osqp.c:1012-1279, step 1: overwrite the raw (un-scaled) values, all of
// them or by CSC index list.  The host then re-runs osqp_cuda_engine_scale (the
// reference un-scales, overwrites and re-scales) and osqp_cuda_engine_refactor.
extern "C" c_int osqp_cuda_engine_update_matrices(osqp_cuda_engine *e, const c_float *Px_new,
                                                  const c_int *Px_idx, c_int P_n,
                                                  const c_float *Ax_new, const c_int *Ax_idx,
                                                  c_int A_n) {
  if (Px_new && mats_update_raw(e->W, true, Px_new, Px_idx, P_n, e->st)) return 1;
  if (Ax_new && mats_update_raw(e->W, false, Ax_new, Ax_idx, A_n, e->st)) return 1;
  return 0;
}
// step 2 (replaces the QDLDL re-factorisation, qdldl_interface.c:381-393): new Jacobi
// diagonal + curvature probe.  0, or -1 when the new K is not positive definite.
extern "C" c_int osqp_cuda_engine_refactor(osqp_cuda_engine *e) {
  if (!e->linsys_ready) return 0;
  const c_int rc = engine_refresh_linsys(e, true);
  if (rc == OSQP_NONCVX_ERROR) return -1;
  return rc;
}

// reference src/auxil.c:524-562 + src/scaling.c:177-192 (un-scale fused with the copy-out)
extern "C" c_int osqp_cuda_engine_get_solution(osqp_cuda_engine *e, c_float *x, c_float *y) {
  const int n = e->n, m = e->m;
  if (x) {
    k_mul<<<EW(n)>>>(e->xt, e->x, e->D, n);
    d2h(x, e->xt, n, e->st);
  }
  if (y && m) {
    k_mul_scal<<<EW(m)>>>(e->pcg.w, e->y, e->E, e->cs + 1, m);
    d2h(y, e->pcg.w, m, e->st);
  }
  CUDA_TRY(cudaStreamSynchronize(e->st));
  return 0;
}
// Solutions of many engines (lockstep batches): the un-scaling kernels and copies of all engines
// are enqueued into pinned staging without a synchronisation per problem; one device-wide wait
// per group of problems, then plain host copies into the callers' (pageable) arrays.
extern "C" c_int osqp_cuda_batch_get_solutions(osqp_cuda_engine **es, c_int count, c_float **xs,
                                               c_float **ys) {
  static double *stage = nullptr;          // guarded by g_bs.mu, like the other batch staging
  static size_t stage_len = 0;
  const size_t GROUP_DOUBLES = (size_t)4 << 20;      // 32 MB of pinned staging
  pthread_mutex_lock(&g_bs.mu);
  int rc = 0;
  c_int i0 = 0;
  while (i0 < count && !rc) {
    size_t need = 0;
    c_int i1 = i0;
    while (i1 < count && (i1 == i0 || need + (size_t)es[i1]->n + es[i1]->m <= GROUP_DOUBLES)) {
      need += (size_t)es[i1]->n + es[i1]->m;
      ++i1;
    }
    if (need > stage_len) {
      if (stage) cudaFreeHost(stage);
      stage = nullptr; stage_len = 0;
      const size_t len = need > GROUP_DOUBLES ? need : GROUP_DOUBLES;
      if (cudaMallocHost((void **)&stage, len * sizeof(double)) != cudaSuccess) { rc = 1; break; }
      stage_len = len;
    }
    // one un-scaling launch (a CTA per problem, same arithmetic as k_mul / k_mul_scal) into a device
    // staging buffer and ONE copy per group, instead of two launches and two copies per problem
    static double *dstage = nullptr;
    static size_t dstage_len = 0;
    if (need > dstage_len) {
      if (dstage) cudaFree(dstage);
      dstage = nullptr; dstage_len = 0;
      const size_t len = need > GROUP_DOUBLES ? need : GROUP_DOUBLES;
      if (cudaMalloc((void **)&dstage, len * sizeof(double)) != cudaSuccess) { rc = 1; break; }
      dstage_len = len;
    }
    if ((!g_bs.st && cudaStreamCreateWithFlags(&g_bs.st, cudaStreamNonBlocking) != cudaSuccess) ||
        batch_args_reserve((size_t)(i1 - i0) * sizeof(BatchUnscaleArgs))) { rc = 1; break; }
    BatchUnscaleArgs *ha = (BatchUnscaleArgs *)g_bb_host;
    size_t off = 0;
    for (c_int i = i0; i < i1; ++i) {
      osqp_cuda_engine *e = es[i];
      ha[i - i0] = BatchUnscaleArgs{e->x, e->D, e->y, e->E, e->cs, e->n, e->m, (long long)off};
      off += (size_t)e->n + e->m;
    }
    if (cudaDeviceSynchronize() != cudaSuccess) { rc = 1; break; }       // the engines' own streams are done
    cudaMemcpyAsync(g_bb_dev, ha, (size_t)(i1 - i0) * sizeof(BatchUnscaleArgs), cudaMemcpyHostToDevice, g_bs.st);
    batch_unscale_kernel<<<(int)(i1 - i0), OSQP_CUDA_THREADS, 0, g_bs.st>>>((const BatchUnscaleArgs *)g_bb_dev, dstage);
    cudaMemcpyAsync(stage, dstage, need * sizeof(double), cudaMemcpyDeviceToHost, g_bs.st);
    if (cudaDeviceSynchronize() != cudaSuccess) { rc = 1; break; }
    off = 0;
    for (c_int i = i0; i < i1; ++i) {
      const int n = es[i]->n, m = es[i]->m;
      if (xs[i]) memcpy(xs[i], stage + off, sizeof(double) * n);
      if (ys[i] && m) memcpy(ys[i], stage + off + n, sizeof(double) * m);
      off += (size_t)n + m;
    }
    i0 = i1;
  }
  pthread_mutex_unlock(&g_bs.mu);
  return rc;
}
extern "C" c_int osqp_cuda_engine_get_iterates(osqp_cuda_engine *e, c_float *x, c_float *z, c_float *y) {
  int bad = d2h(x, e->x, e->n, e->st) | d2h(z, e->z, e->m, e->st) | d2h(y, e->y, e->m, e->st);
  CUDA_TRY(cudaStreamSynchronize(e->st));
  return bad;
}
// reference src/auxil.c:754-782: certificates, scaled back by D / E when requested
extern "C" c_int osqp_cuda_engine_get_certificates(osqp_cuda_engine *e, c_float *delta_x,
                                                   c_float *delta_y, c_int unscale) {
  const int n = e->n, m = e->m;
  if (delta_x) {
    if (unscale) { k_mul<<<EW(n)>>>(e->xt, e->dx, e->D, n); d2h(delta_x, e->xt, n, e->st); }
    else d2h(delta_x, e->dx, n, e->st);
  }
  if (delta_y && m) {
    if (unscale) { k_mul<<<EW(m)>>>(e->pcg.w, e->dyp, e->E, m); d2h(delta_y, e->pcg.w, m, e->st); }
    else d2h(delta_y, e->dyp, m, e->st);
  }
  CUDA_TRY(cudaStreamSynchronize(e->st));
  return 0;
}

extern "C" c_int osqp_cuda_engine_get_stats(const osqp_cuda_engine *e, OSQPB200Stats *out) {
  *out = e->stats;
  out->kernel_launches = e->L.launches;
  out->sell_launches = e->L.sell_launches;
  out->tile_launches = e->L.tile_launches;
  out->slack_vars = e->slack.on ? e->slack.ny : 0;
  return 0;
}
extern "C" c_int osqp_cuda_engine_profile(osqp_cuda_engine *e, c_int enable) {
  e->L.collect();
  e->L.profiling = enable != 0;
  if (enable) for (int k = 0; k < OSQP_CUDA_K_COUNT; ++k) { e->L.ms[k] = 0; e->L.cnt[k] = 0; }
  return 0;
}
extern "C" c_int osqp_cuda_engine_profile_get(osqp_cuda_engine *e, c_int kernel_id, c_int *launches,
                                              c_float *total_ms, c_float *bytes_per_launch) {
  if (kernel_id < 0 || kernel_id >= OSQP_CUDA_K_COUNT) return 1;
  e->L.collect();
  if (launches) *launches = e->L.cnt[kernel_id];
  if (total_ms) *total_ms = e->L.ms[kernel_id];
  if (bytes_per_launch) *bytes_per_launch = e->bytes_per_launch[kernel_id];
  return 0;
}
extern "C" c_int osqp_cuda_engine_sync(osqp_cuda_engine *e) {
  CUDA_TRY(cudaStreamSynchronize(e->st));
  CUDA_TRY(cudaGetLastError());
  return 0;
}

extern "C" c_int osqp_cuda_engine_timer_start(osqp_cuda_engine *e) {
  if (!e->ev_a) { CUDA_TRY(cudaEventCreate(&e->ev_a)); CUDA_TRY(cudaEventCreate(&e->ev_b)); }
  CUDA_TRY(cudaEventRecord(e->ev_a, e->st));
  return 0;
}
extern "C" c_int osqp_cuda_engine_timer_stop(osqp_cuda_engine *e, c_float *elapsed_ms) {
  float t = 0.f;
  if (!e->ev_a) return 1;
  CUDA_TRY(cudaEventRecord(e->ev_b, e->st));
  CUDA_TRY(cudaEventSynchronize(e->ev_b));
  CUDA_TRY(cudaEventElapsedTime(&t, e->ev_a, e->ev_b));
  if (elapsed_ms) *elapsed_ms = t;
  return 0;
}

// ==================================================================== polish
// K12 -- reference src/polish.c:19-350.  The reduced KKT system
//   [P + dI, Ared' ; Ared, -dI] [x ; nu] = [-q ; b_red]
// is solved in its reduced form (P + dI + Ared' Ared / d) x = -q + Ared' b_red / d by the
// same CG; "Ared" is never built: inactive rows get rho = 0, active rows rho = 1/d.
__global__ void k_polish_active(const double *z, const double *y, const double *l, const double *u,
                                double delta, int m, int *act, double *rho, double *b, double *t,
                                int *counts) {
  EW_LOOP(i, m) {
    int a = 0;
    if (z[i] - l[i] < -y[i]) a = 1;            // lower-active (polish.c:34)
    else if (u[i] - z[i] < y[i]) a = 2;        // upper-active (polish.c:43)
    act[i] = a;
    const double r = a ? 1.0 / delta : 0.0;
    const double bi = a == 1 ? l[i] : (a == 2 ? u[i] : 0.0);
    rho[i] = r; b[i] = bi; t[i] = r * bi;
    if (a == 1) atomicAdd(counts + 1, 1);
    if (a == 2) atomicAdd(counts + 2, 1);
  }
}
struct EpiPolRhs {       // rhs = -q + A' t       (split sums: only the A' part is used)
  static constexpr int NSUM = 0, NMAX = 0;
  const double *q; double *rhs;
  __device__ bool begin(double *) { return true; }
  EPI_NO_PREFETCH
  __device__ void row(int j, const double *s, const Pre &, double *, double *) const { rhs[j] = -q[j] + s[1]; }
};
struct EpiPolNu {        // nu = rho .* (A x - b)
  static constexpr int NSUM = 0, NMAX = 0;
  const double *rho, *b; double *nu;
  __device__ bool begin(double *) { return true; }
  EPI_NO_PREFETCH
  __device__ void row(int i, const double *s, const Pre &, double *, double *) const { nu[i] = rho[i] * (s[0] - b[i]); }
};
struct EpiPolR2 {        // t = rho .* (b - A x) ; g = nu - t      (refinement, polish.c:150-167)
  static constexpr int NSUM = 0, NMAX = 0;
  const double *rho, *b, *nu; double *t, *g;
  __device__ bool begin(double *) { return true; }
  EPI_NO_PREFETCH
  __device__ void row(int i, const double *s, const Pre &, double *, double *) const {
    const double ti = rho[i] * (b[i] - s[0]);
    t[i] = ti; g[i] = nu[i] - ti;
  }
};
struct EpiPolRefRhs {    // rhs = -q - (P x + A' g)
  static constexpr int NSUM = 0, NMAX = 0;
  const double *q; double *rhs;
  __device__ bool begin(double *) { return true; }
  EPI_NO_PREFETCH
  __device__ void row(int j, const double *s, const Pre &, double *, double *) const { rhs[j] = -q[j] - s[0]; }
};
struct EpiPolDnu {       // nu += rho .* (A dx) - t
  static constexpr int NSUM = 0, NMAX = 0;
  const double *rho, *t; double *nu;
  __device__ bool begin(double *) { return true; }
  EPI_NO_PREFETCH
  __device__ void row(int i, const double *s, const Pre &, double *, double *) const { nu[i] += rho[i] * s[0] - t[i]; }
};
template <> struct epi_tile_ok<EpiPolNu> { static constexpr bool value = true; };
template <> struct epi_tile_ok<EpiPolR2> { static constexpr bool value = true; };
template <> struct epi_tile_ok<EpiPolRefRhs> { static constexpr bool value = true; };
template <> struct epi_tile_ok<EpiPolDnu> { static constexpr bool value = true; };
__global__ void k_add(double *a, const double *b, int len) { EW_LOOP(i, len) a[i] += b[i]; }
// reference src/proj.c:16-29
__global__ void k_normalcone(double *z, double *y, const double *l, const double *u, int m) {
  EW_LOOP(i, m) {
    const double t = z[i] + y[i];
    const double zn = fmin(fmax(t, l[i]), u[i]);
    z[i] = zn; y[i] = t - zn;
  }
}

extern "C" c_int osqp_cuda_engine_polish(osqp_cuda_engine *e, c_float delta, c_int refine_iter,
                                         c_int *n_low, c_int *n_upp) {
  const int n = e->n, m = e->m;
  const size_t N = n, M = m ? m : 1;
  PcgDev &S = e->pcg;
  if (!e->pol_x) {
    if (dev_alloc(&e->pol_x, N) || dev_alloc(&e->pol_rhs, N) || dev_alloc(&e->pol_z, M) ||
        dev_alloc(&e->pol_y, M) || dev_alloc(&e->pol_rho, M) || dev_alloc(&e->pol_b, M) ||
        dev_alloc(&e->pol_nu, M) || dev_alloc(&e->pol_t, M) || dev_alloc(&e->pol_act, M)) return 1;
  }
  int counts[4] = {0, 0, 0, 0};
  cudaMemsetAsync(e->flag, 0, sizeof(int) * 4, e->st);
  if (m) k_polish_active<<<EW(m)>>>(e->z, e->y, e->l, e->u, delta, m, e->pol_act, e->pol_rho,
                                    e->pol_b, e->pol_t, e->flag);
  CUDA_TRY(cudaMemcpyAsync(counts, e->flag, sizeof(counts), cudaMemcpyDeviceToHost, e->st));
  // the polish system: rho <- pol_rho, sigma <- delta, new Jacobi diagonal
  S.rho = e->pol_rho; S.sigma = delta;
  pcg_jacobi(S, e->L);
  EpiPolRhs er{e->q, e->pol_rhs};
  launch_stream<1, true, 0>(e->L, -1, e->W.M, S.gridM, Gather{e->x, e->pol_t}, Gather{nullptr, nullptr},
                            er, nullptr);
  cudaMemcpyAsync(e->pol_x, e->x, sizeof(double) * N, cudaMemcpyDeviceToDevice, e->st);
  const int cap = n < 500 ? 10 * n + 50 : 5000;
  int conv = 0, rc = 0;
  if (pcg_solve_sync(S, e->L, e->pol_rhs, e->pol_x, false, 1e-10, cap, &conv) < 0) rc = 1;
  if (!rc && m) {
    EpiPolNu en{e->pol_rho, e->pol_b, e->pol_nu};
    launch_stream<1, false, 0>(e->L, -1, e->W.A, S.gridA, Gather{e->pol_x, nullptr},
                               Gather{nullptr, nullptr}, en, nullptr);
  }
  // iterative refinement against the UN-regularised reduced KKT (polish.c:134-181)
  for (int it = 0; !rc && it < (int)refine_iter; ++it) {
    if (m) {
      EpiPolR2 e2{e->pol_rho, e->pol_b, e->pol_nu, e->pol_t, e->pol_z};
      launch_stream<1, false, 0>(e->L, -1, e->W.A, S.gridA, Gather{e->pol_x, nullptr},
                                 Gather{nullptr, nullptr}, e2, nullptr);
    }
    EpiPolRefRhs e3{e->q, e->pol_rhs};
    launch_stream<1, false, 0>(e->L, -1, e->W.M, S.gridM, Gather{e->pol_x, e->pol_z},
                               Gather{nullptr, nullptr}, e3, nullptr);
    if (pcg_solve_sync(S, e->L, e->pol_rhs, e->xt, true, 1e-10, cap, &conv) < 0) { rc = 1; break; }
    if (m) {
      EpiPolDnu e4{e->pol_rho, e->pol_t, e->pol_nu};
      launch_stream<1, false, 0>(e->L, -1, e->W.A, S.gridA, Gather{e->xt, nullptr},
                                 Gather{nullptr, nullptr}, e4, nullptr);
    }
    k_add<<<EW(n)>>>(e->pol_x, e->xt, n);
  }
  if (!rc && m) {
    EpiStore es{e->pol_z};
    launch_stream<1, false, 0>(e->L, -1, e->W.A, S.gridA, Gather{e->pol_x, nullptr},
                               Gather{nullptr, nullptr}, es, nullptr);
    cudaMemcpyAsync(e->pol_y, e->pol_nu, sizeof(double) * M, cudaMemcpyDeviceToDevice, e->st);
    k_normalcone<<<EW(m)>>>(e->pol_z, e->pol_y, e->l, e->u, m);
  }
  // back to the ADMM system
  S.rho = e->rho_vec; S.sigma = e->sigma;
  engine_refresh_linsys(e, false);
  CUDA_TRY(cudaStreamSynchronize(e->st));
  CUDA_TRY(cudaGetLastError());
  if (n_low) *n_low = counts[1];
  if (n_upp) *n_upp = counts[2];
  return rc;
}

// reference src/polish.c:323-327: the polished point becomes the iterate
extern "C" c_int osqp_cuda_engine_polish_accept(osqp_cuda_engine *e) {
  const size_t N = e->n, M = e->m;
  cudaMemcpyAsync(e->x, e->pol_x, sizeof(double) * N, cudaMemcpyDeviceToDevice, e->st);
  if (M) {
    cudaMemcpyAsync(e->z, e->pol_z, sizeof(double) * M, cudaMemcpyDeviceToDevice, e->st);
    cudaMemcpyAsync(e->y, e->pol_y, sizeof(double) * M, cudaMemcpyDeviceToDevice, e->st);
  }
  e->v_dirty = true;
  CUDA_TRY(cudaStreamSynchronize(e->st));
  return 0;
}

// =================================================================== plugin
// The reference's linsys_solver contract with HOST pointers (strict mode).  An
// instance either owns its device matrices (init_linsys_solver_cuda_pcg) or is the
// view of an engine's linear system (work->linsys_solver in osqp_setup).
struct PluginImpl {
  bool owns = true;
  osqp_cuda_engine *engine = nullptr;   // when !owns
  DevMats *W = nullptr;
  PcgDev *pcg = nullptr;
  Launcher *L = nullptr;
  cudaStream_t st = 0;
  double *rho = nullptr;        // m (owned only when owns)
  double *bdev = nullptr;       // n + m
  double *tdev = nullptr;       // m
  double *rhs = nullptr;        // n
  double *xsol = nullptr;       // n, warm start = previous solution
  double *tdev_scalar = nullptr; // 1: ||rhs||inf of the current solve
  bool have_x = false;
  double eps_rel = 1e-12;
  int max_iter = 0;
  int last_iters = 0;
};

struct EpiPlugRhs {      // rhs = b1 + A' t   (split sums: A' part only)
  static constexpr int NSUM = 0, NMAX = 0;
  const double *b1; double *rhs;
  __device__ bool begin(double *) { return true; }
  EPI_NO_PREFETCH
  __device__ void row(int j, const double *s, const Pre &, double *, double *) const { rhs[j] = b1[j] + s[1]; }
};

static int plugin_alloc_buffers(PluginImpl *I) {
  const size_t N = I->W->n, M = I->W->m;
  if (dev_alloc(&I->bdev, N + M) || dev_alloc(&I->tdev, M) || dev_alloc(&I->rhs, N) ||
      dev_alloc(&I->xsol, N) || dev_alloc(&I->tdev_scalar, 1)) return 1;
  cudaMemsetAsync(I->xsol, 0, sizeof(double) * N, I->st);
  I->max_iter = (int)(N < 500 ? 10 * N + 50 : 5000);
  return 0;
}

// ||v||inf of a device vector into *out (device; zeroed by the caller)
__global__ void k_absmax(const double *v, int len, double *out) {
  double a = 0.0;
  EW_LOOP(i, len) a = fmax(a, fabs(v[i]));
  a = warp_max(a);
  if ((threadIdx.x & 31) == 0 && a > 0.0) atomic_max_nonneg(out, a);
}

extern "C" c_int solve_linsys_cuda_pcg(cuda_pcg_solver *self, c_float *b) {
  PluginImpl *I = (PluginImpl *)self->impl;
  PcgDev &S = *I->pcg;
  const int n = I->W->n, m = I->W->m;
  cudaStream_t st = I->st;
  if (!I->owns) { S.rho = I->engine->rho_vec; S.sigma = I->engine->sigma; }
  const double *rho = S.rho;
  CUDA_TRY(cudaMemcpyAsync(I->bdev, b, sizeof(double) * (n + m), cudaMemcpyHostToDevice, st));
  if (m) k_mul<<<elementwise_grid(m), OSQP_CUDA_THREADS, 0, st>>>(I->tdev, rho, I->bdev + n, m);
  EpiPlugRhs er{I->bdev, I->rhs};
  launch_stream<1, true, 0>(*I->L, -1, I->W->M, S.gridM, Gather{I->bdev, I->tdev},
                            Gather{nullptr, nullptr}, er, nullptr);
  // The target is relative to ||rhs||inf, not to the residual of the warm start: as ADMM converges the
  // warm start gets better, and a target relative to ITS residual would sink below the round-off of
  // K x, send the CG to its iteration cap and make this call report failure for nothing.
  double rhs_norm = 0.0;
  cudaMemsetAsync(I->tdev_scalar, 0, sizeof(double), st);
  k_absmax<<<grid_for(n), OSQP_CUDA_THREADS, 0, st>>>(I->rhs, n, I->tdev_scalar);
  CUDA_TRY(cudaMemcpyAsync(&rhs_norm, I->tdev_scalar, sizeof(double), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  const double save_floor = S.abs_floor;
  S.abs_floor = fmax(save_floor, I->eps_rel * rhs_norm);
  int conv = 0;
  const int iters = pcg_solve_sync(S, *I->L, I->rhs, I->xsol, !I->have_x, I->eps_rel, I->max_iter, &conv);
  S.abs_floor = save_floor;
  if (iters < 0) return 1;
  I->have_x = true;
  I->last_iters = iters;
  if (m) {
    if (self->polish) {     // nu = (A x - b2) / delta          (qdldl_interface.c:354-356)
      EpiPolNu en{rho, I->bdev + n, I->tdev};
      launch_stream<1, false, 0>(*I->L, -1, I->W->A, S.gridA, Gather{I->xsol, nullptr},
                                 Gather{nullptr, nullptr}, en, nullptr);
    } else {                // z~ = A x~                        (qdldl_interface.c:368-370)
      EpiStore es{I->tdev};
      launch_stream<1, false, 0>(*I->L, -1, I->W->A, S.gridA, Gather{I->xsol, nullptr},
                                 Gather{nullptr, nullptr}, es, nullptr);
    }
    CUDA_TRY(cudaMemcpyAsync(b + n, I->tdev, sizeof(double) * m, cudaMemcpyDeviceToHost, st));
  }
  CUDA_TRY(cudaMemcpyAsync(b, I->xsol, sizeof(double) * n, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  PcgScal h;
  CUDA_TRY(cudaMemcpy(&h, S.S, sizeof(PcgScal), cudaMemcpyDeviceToHost));
  if (h.noncvx) return 2;
  return conv ? 0 : 3;
}

static c_int plugin_probe(PluginImpl *I) {
  PcgDev &S = *I->pcg;
  const int n = I->W->n;
  std::vector<double> b(n);
  unsigned long long s = 0x9E3779B97F4A7C15ull;
  for (int i = 0; i < n; ++i) {
    s = s * 6364136223846793005ull + 1442695040888963407ull;
    b[i] = ((double)(s >> 11) / 9007199254740992.0) - 0.5;
  }
  cudaMemcpyAsync(I->rhs, b.data(), sizeof(double) * n, cudaMemcpyHostToDevice, I->st);
  cudaMemsetAsync(&S.S->noncvx, 0, sizeof(int), I->st);
  pcg_jacobi(S, *I->L);
  int conv = 0;
  const int it = pcg_solve_sync(S, *I->L, I->rhs, I->bdev, true, 1e-12, n < 30 ? n + 2 : 30, &conv);
  PcgScal h;
  if (cudaMemcpy(&h, S.S, sizeof(PcgScal), cudaMemcpyDeviceToHost) != cudaSuccess || it < 0)
    return OSQP_LINSYS_SOLVER_INIT_ERROR;
  return h.noncvx ? OSQP_NONCVX_ERROR : 0;
}

extern "C" c_int update_linsys_solver_rho_vec_cuda_pcg(cuda_pcg_solver *self, const c_float *rho_vec) {
  PluginImpl *I = (PluginImpl *)self->impl;
  const int m = I->W->m;
  double *dst = I->owns ? I->rho : I->engine->rho_vec;
  if (m) CUDA_TRY(cudaMemcpyAsync(dst, rho_vec, sizeof(double) * m, cudaMemcpyHostToDevice, I->st));
  if (!I->owns) return engine_refresh_linsys(I->engine, false);
  pcg_jacobi(*I->pcg, *I->L);
  CUDA_TRY(cudaStreamSynchronize(I->st));
  return 0;
}

extern "C" c_int update_linsys_solver_matrices_cuda_pcg(cuda_pcg_solver *self, const csc *P, const csc *A) {
  PluginImpl *I = (PluginImpl *)self->impl;
  if (!I->owns) {
    fprintf(stderr, "osqp_cuda_pcg: this instance is the view of an engine; use osqp_update_P/A\n");
    return 1;
  }
  // same sparsity, new (already scaled) values -- reference update_KKT_P / update_KKT_A
  if (mats_update_raw(*I->W, true, P->x, nullptr, 0, I->st) ||
      mats_update_raw(*I->W, false, A->x, nullptr, 0, I->st) || mats_load_raw(*I->W, I->st)) return 1;
  mats_sell_sync(*I->W, I->st);
  I->have_x = false;
  const c_int rc = plugin_probe(I);
  return rc == OSQP_NONCVX_ERROR ? -1 : rc;
}

extern "C" void free_linsys_solver_cuda_pcg(cuda_pcg_solver *self) {
  if (!self) return;
  PluginImpl *I = (PluginImpl *)self->impl;
  if (I) {
    if (I->st) cudaStreamSynchronize(I->st);
    dev_free(I->bdev); dev_free(I->tdev); dev_free(I->rhs); dev_free(I->xsol); dev_free(I->tdev_scalar);
    if (I->owns) {
      dev_free(I->rho);
      if (I->L) { I->L->destroy(); delete I->L; }
      if (I->pcg) { pcg_release(*I->pcg); delete I->pcg; }
      if (I->W) { mats_free(*I->W); delete I->W; }
      if (I->st) cudaStreamDestroy(I->st);
      delete I;
      free(self);
    } else {
      // the engine owns both the view struct and the device state; osqp_cleanup frees
      // the engine afterwards, so only the staging buffers go here
      I->engine->solver = nullptr;
      delete I;
      free(self);
    }
  }
}

static void plugin_fill_vtable(cuda_pcg_solver *s, int n, int m, int polish, PluginImpl *I) {
  s->type = CUDA_PCG_SOLVER;
  s->solve = &solve_linsys_cuda_pcg;
  s->free = &free_linsys_solver_cuda_pcg;
  s->update_matrices = &update_linsys_solver_matrices_cuda_pcg;
  s->update_rho_vec = &update_linsys_solver_rho_vec_cuda_pcg;
  s->nthreads = 1;
  s->n = n; s->m = m; s->polish = polish;
  s->impl = I;
}

static void plugin_bind_engine(osqp_cuda_engine *e) {
  if (e->solver) return;
  cuda_pcg_solver *s = (cuda_pcg_solver *)calloc(1, sizeof(cuda_pcg_solver));
  PluginImpl *I = new PluginImpl();
  I->owns = false; I->engine = e; I->W = &e->W; I->pcg = &e->pcg; I->L = &e->L; I->st = e->st;
  if (!s || plugin_alloc_buffers(I)) { delete I; free(s); return; }
  plugin_fill_vtable(s, e->n, e->m, 0, I);
  e->solver = s;
}

extern "C" c_int init_linsys_solver_cuda_pcg(cuda_pcg_solver **sp, const csc *P, const csc *A,
                                             c_float sigma, const c_float *rho_vec, c_int polish) {
  *sp = nullptr;
  if (bind_device()) return OSQP_LINSYS_SOLVER_INIT_ERROR;
  cuda_pcg_solver *s = (cuda_pcg_solver *)calloc(1, sizeof(cuda_pcg_solver));
  PluginImpl *I = new PluginImpl();
  I->W = new DevMats(); I->pcg = new PcgDev(); I->L = new Launcher();
  plugin_fill_vtable(s, (int)P->n, (int)A->m, (int)polish, I);
  c_int rc = OSQP_LINSYS_SOLVER_INIT_ERROR;
  do {
    if (cudaStreamCreateWithFlags(&I->st, cudaStreamNonBlocking) != cudaSuccess) break;
    I->L->st = I->st;
    if (mats_build(*I->W, P, A, I->st)) break;
    if (pcg_alloc(*I->pcg, I->W, I->st)) break;
    const int m = I->W->m;
    if (dev_alloc(&I->rho, (size_t)m) || plugin_alloc_buffers(I)) break;
    std::vector<double> r(m ? m : 1);
    for (int i = 0; i < m; ++i) r[i] = polish ? 1.0 / sigma : rho_vec[i];
    if (m && cudaMemcpyAsync(I->rho, r.data(), sizeof(double) * m, cudaMemcpyHostToDevice, I->st)) break;
    if (cudaStreamSynchronize(I->st) != cudaSuccess) break;
    I->pcg->rho = I->rho; I->pcg->sigma = sigma;
    rc = plugin_probe(I);
  } while (0);
  if (rc) { free_linsys_solver_cuda_pcg(s); return rc; }
  *sp = s;
  return 0;
}

extern "C" void cuda_pcg_set_tolerance(cuda_pcg_solver *self, c_float eps_rel, c_int max_iter) {
  PluginImpl *I = (PluginImpl *)self->impl;
  if (eps_rel > 0) I->eps_rel = eps_rel;
  if (max_iter > 0) I->max_iter = (int)max_iter;
}
extern "C" c_int cuda_pcg_uses_tile(const cuda_pcg_solver *self) {
  const PluginImpl *I = (const PluginImpl *)self->impl;
  return (I->W->A.tile.ok ? 1 : 0) | (I->W->M.tile.ok ? 2 : 0);
}
extern "C" c_int cuda_pcg_last_iters(const cuda_pcg_solver *self) {
  return ((const PluginImpl *)self->impl)->last_iters;
}

// ---- kernel-level check points (tests only): the SpMV launches of the solve path on caller data.
// which = 0: y (m) = A x (n)                            -- K_SPMV_A's product, reference mat_vec, src/lin_alg.c:241-271
// which = 1: y (n) = [Pf | A'] [x (n) ; x + n (m)]      -- K_SPMV_M's product, reference mat_tpose_vec :273-322 with
//                                                          the symmetric-P completion of src/auxil.c:296-306
// which = 2: y (2 m) = [A x0 | A x1], x = [x0 (n) ; x1 (n)]   -- the two-vector pass of the residual kernels (NV = 2)
// The matrices are the engine's current (scaled) device copies; the launch takes whatever kernel the
// solve path would take (column-blocked mirror, sliced-ELL mirror or CSR).  HOST pointers.
struct EpiStore2 {
  static constexpr int NSUM = 0, NMAX = 0;
  double *out; int stride;
  __device__ bool begin(double *) { return true; }
  EPI_NO_PREFETCH
  __device__ void row(int i, const double *s, const Pre &, double *, double *) const { out[i] = s[0]; out[stride + i] = s[1]; }
};
extern "C" c_int osqp_cuda_engine_spmv(osqp_cuda_engine *e, c_int which, const c_float *x, c_float *y) {
  if (!e || which < 0 || which > 2) return 1;
  const size_t n = e->n, m = e->m;
  const size_t nin = which == 0 ? n : (which == 1 ? n + m : 2 * n), nout = which == 0 ? m : (which == 1 ? n : 2 * m);
  if ((which != 1 && m == 0) || !nout) return 0;
  double *din = nullptr, *dhi = nullptr, *dout = nullptr;
  int rc = 1;
  do {
    // the m-part of the gathered vector gets its own (aligned) allocation, as on the solve path
    if (dev_alloc(&din, which == 1 ? n : nin) || dev_alloc(&dhi, which == 0 ? 1 : (which == 1 ? m : n)) ||
        dev_alloc(&dout, nout)) break;
    if (cudaMemcpyAsync(din, x, sizeof(double) * n, cudaMemcpyHostToDevice, e->st) != cudaSuccess) break;
    if (which == 1 && m && cudaMemcpyAsync(dhi, x + n, sizeof(double) * m, cudaMemcpyHostToDevice, e->st) != cudaSuccess) break;
    if (which == 2 && cudaMemcpyAsync(dhi, x + n, sizeof(double) * n, cudaMemcpyHostToDevice, e->st) != cudaSuccess) break;
    if (which == 0) {
      EpiStore es{dout};
      launch_stream<1, false, 0>(e->L, -1, e->W.A, e->pcg.gridA, Gather{din, nullptr}, Gather{nullptr, nullptr}, es, nullptr);
    } else if (which == 1) {
      EpiStore es{dout};
      launch_stream<1, false, 0>(e->L, -1, e->W.M, e->pcg.gridM, Gather{din, dhi}, Gather{nullptr, nullptr}, es, nullptr);
    } else {
      EpiStore2 es{dout, (int)m};
      launch_stream<2, false, 0>(e->L, -1, e->W.A, e->pcg.gridA, Gather{din, nullptr}, Gather{dhi, nullptr}, es, nullptr);
    }
    if (cudaMemcpyAsync(y, dout, sizeof(double) * nout, cudaMemcpyDeviceToHost, e->st) != cudaSuccess) break;
    if (cudaStreamSynchronize(e->st) != cudaSuccess || cudaGetLastError() != cudaSuccess) break;
    rc = 0;
  } while (0);
  if (rc) fprintf(stderr, "osqp_cuda_pcg: osqp_cuda_engine_spmv failed: %s\n", cudaGetErrorString(cudaGetLastError()));
  dev_free(din); dev_free(dhi); dev_free(dout);
  return rc;
}

// ====================================================== allocation-free solution read-back
// osqp_cuda_engine_get_solution_noalloc: the un-scaling of osqp_cuda_engine_get_solution
// (reference src/auxil.c:524-562 + src/scaling.c:177-192) for callers that may not allocate
// device memory.  No counterpart in the reference.
// The staging area that osqp_cuda_engine_get_solution takes from the engine's cudaMalloc'ed
// vectors (xt, pcg.w) is a file-scope device array here: OSQP_CUDA_NOALLOC_MAX doubles (2 MB),
// held by every GPU that loads this library until the process exits.  Every call shares it and
// it keeps the last call's contents: each call overwrites the n + m slots it reads back (nothing
// has to start at zero), and calls that overlap on the GPU race on it.
static __device__ __align__(256) double g_noalloc_scratch[OSQP_CUDA_NOALLOC_MAX];

// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address
// the kernels get is resolved once per device, with that device current.
static double *noalloc_scratch_ptr(int device) {
  static std::mutex mu;
  static double *ptr[64] = {nullptr};
  if (device < 0 || device >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[device]) {
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    void *p = nullptr;
    const cudaError_t err = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (err != cudaSuccess || !p) {
      fprintf(stderr, "osqp_cuda_pcg: cudaGetSymbolAddress(g_noalloc_scratch) failed on device %d: %s\n",
              device, cudaGetErrorString(err));
      cudaGetLastError();
      return nullptr;
    }
    ptr[device] = (double *)p;
  }
  return ptr[device];
}

extern "C" c_int osqp_cuda_noalloc_max(void) { return (c_int)OSQP_CUDA_NOALLOC_MAX; }

extern "C" c_int osqp_cuda_engine_get_solution_noalloc(osqp_cuda_engine *e, c_float *x, c_float *y) {
  if (!e) return 1;
  const long long n = e->n, m = e->m;
  if (n + m > (long long)OSQP_CUDA_NOALLOC_MAX) {
    fprintf(stderr, "osqp_cuda_pcg: osqp_cuda_engine_get_solution_noalloc: n + m = %lld exceeds the static "
            "scratch of %d doubles; use osqp_cuda_engine_get_solution\n", n + m, (int)OSQP_CUDA_NOALLOC_MAX);
    return 2;
  }
  double *scratch = noalloc_scratch_ptr(g_device);
  if (!scratch) return 1;
  if (x) {
    k_mul<<<EW((int)n)>>>(scratch, e->x, e->D, (int)n);
    d2h(x, scratch, (size_t)n, e->st);
  }
  if (y && m) {
    k_mul_scal<<<EW((int)m)>>>(scratch + n, e->y, e->E, e->cs + 1, (int)m);
    d2h(y, scratch + n, (size_t)m, e->st);
  }
  CUDA_TRY(cudaStreamSynchronize(e->st));
  CUDA_TRY(cudaGetLastError());
  return 0;
}
