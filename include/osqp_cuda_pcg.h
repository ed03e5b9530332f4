/*
 * osqp_cuda_pcg.h -- C-ABI of libosqp_cuda_pcg.so (hand-written CUDA, sm_100a).
 *
 * Plain pointers and sizes only; no torch / C++ types cross this boundary.
 * Two groups of entry points:
 *
 *  (1) The reference's linear-system PLUGIN contract, strict host-pointer form:
 *      the five functions a backend provides (reference
 *      lin_sys/direct/qdldl/qdldl_interface.h:89-128,
 *      lin_sys/direct/pardiso/pardiso_interface.h:86-121), reached through a
 *      struct whose head is layout-compatible with struct linsys_solver
 *      (reference include/types.h:298-319).
 *
 *  (2) The device-resident ADMM ENGINE used by this repo's osqp_solve(): it
 *      owns P, A, A' and every iterate in HBM, so only a few scalars cross to
 *      the host per termination check.  These replace the host loops of
 *      reference src/auxil.c, src/lin_alg.c, src/scaling.c, src/proj.c and
 *      src/polish.c cited at each declaration.
 */
#ifndef OSQP_CUDA_PCG_H
#define OSQP_CUDA_PCG_H

#include "osqp.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ======================================================= (1) plugin contract */

typedef struct cuda_pcg cuda_pcg_solver;

struct cuda_pcg {
  /* ---- head: identical to struct linsys_solver ---- */
  enum linsys_solver_type type;
  c_int (*solve)(cuda_pcg_solver *self, c_float *b);
  void  (*free)(cuda_pcg_solver *self);
  c_int (*update_matrices)(cuda_pcg_solver *self, const csc *P, const csc *A);
  c_int (*update_rho_vec)(cuda_pcg_solver *self, const c_float *rho_vec);
  c_int nthreads;
  /* ---- private ---- */
  c_int n, m, polish;
  void *impl;
};

/* replaces init_linsys_solver_qdldl (reference qdldl_interface.c:177-313).
 * P upper-triangular CSC n x n, A CSC m x n, both HOST, borrowed for the call.
 * polish != 0: rho_vec is NULL and sigma carries delta (reference
 * src/polish.c:232-234).  Returns 0, OSQP_LINSYS_SOLVER_INIT_ERROR or
 * OSQP_NONCVX_ERROR (K = P + sigma I + A' diag(rho) A found not positive
 * definite by the CG curvature probe).                                        */
c_int init_linsys_solver_cuda_pcg(cuda_pcg_solver **sp, const csc *P, const csc *A,
                                  c_float sigma, const c_float *rho_vec, c_int polish);

/* replaces solve_linsys_qdldl (qdldl_interface.c:350-376).  b: HOST, n+m, in/out.
 * ADMM mode in [sigma x_prev - q ; z_prev - y/rho], out [x~ ; z~ = A x~].
 * Polish mode in rhs, out [x ; nu].                                           */
c_int solve_linsys_cuda_pcg(cuda_pcg_solver *self, c_float *b);
/* replaces update_linsys_solver_matrices_qdldl (qdldl_interface.c:381-393)    */
c_int update_linsys_solver_matrices_cuda_pcg(cuda_pcg_solver *self, const csc *P, const csc *A);
/* replaces update_linsys_solver_rho_vec_qdldl (qdldl_interface.c:396-410)     */
c_int update_linsys_solver_rho_vec_cuda_pcg(cuda_pcg_solver *self, const c_float *rho_vec);
/* replaces free_linsys_solver_qdldl (qdldl_interface.c:22-50)                 */
void  free_linsys_solver_cuda_pcg(cuda_pcg_solver *self);

/* Strict-mode PCG controls (additive): relative residual target and cap.      */
void  cuda_pcg_set_tolerance(cuda_pcg_solver *self, c_float eps_rel, c_int max_iter);
c_int cuda_pcg_last_iters(const cuda_pcg_solver *self);
/* bit 0: SpMVs over A, bit 1: SpMVs over [P | A'] of this instance go through a column-blocked
 * mirror (spmv_tile.cuh)                                                       */
c_int cuda_pcg_uses_tile(const cuda_pcg_solver *self);

/* ============================================================== (2) engine */

typedef struct osqp_cuda_engine osqp_cuda_engine;

/* Scalars produced by one residual pass (K9 + K10); everything osqp_solve's
 * host side needs for update_info / check_termination / adapt_rho
 * (reference src/auxil.c:240-359, 361-512, 13-52, 227-238).
 * "_s" = in the scaled space, "_u" = un-scaled by Einv / Dinv / E / D.        */
typedef struct {
  c_float pri_res_s, pri_res_u;       /* ||Ax - z||inf                         */
  c_float z_norm_s, z_norm_u;         /* ||z||inf                              */
  c_float Ax_norm_s, Ax_norm_u;       /* ||Ax||inf                             */
  c_float dua_res_s, dua_res_u;       /* ||q + Px + A'y||inf (no cinv)         */
  c_float q_norm_s, q_norm_u;
  c_float Aty_norm_s, Aty_norm_u;
  c_float Px_norm_s, Px_norm_u;
  c_float xPx, qx;                    /* objective = cinv (xPx/2 + qx)         */
  c_float dy_norm_s, dy_norm_u;       /* projected delta_y: ||.||, ||E .||     */
  c_float ineq_lhs;                   /* u'(dy)+ + l'(dy)-                     */
  c_float Atdy_norm_s, Atdy_norm_u;   /* ||A'dy||, ||Dinv A'dy||               */
  c_float dx_norm_s, dx_norm_u;       /* ||dx||, ||D dx||                      */
  c_float q_dx;                       /* q'dx                                  */
  c_float Pdx_norm_s, Pdx_norm_u;     /* ||P dx||, ||Dinv P dx||               */
  c_float Adx_maxU_s, Adx_maxU_u;     /* max over rows with finite u of (A dx) */
  c_float Adx_minL_s, Adx_minL_u;     /* min over rows with finite l of (A dx) */
  c_float pcg_noncvx;                 /* != 0: a CG since the last rho / matrix update met p'Kp <= 0 or a
                                       * non-positive Jacobi diagonal: K = P + sigma I + A' rho A is not
                                       * positive definite (what the reference reads off the LDL' inertia,
                                       * qdldl_interface.c:93-98)                                  */
} osqp_cuda_resid;

/* Select the CUDA device for subsequently created engines / solvers.          */
c_int osqp_cuda_set_device(c_int device);
/* 0 if a usable sm_100 device is present, else nonzero (+ message on stderr). */
c_int osqp_cuda_probe(void);

/* ---- Row-partitioned mode (no counterpart in the reference, which is single
 * process).  One process per GPU.  Rank g passes to osqp_setup / osqp_cuda_engine_create its
 * slab of the rows of A, l, u (and gets back its slab of y, z); x, q and every scalar are
 * replicated; P is passed by rank 0 only (the other ranks pass an n x n matrix without
 * entries).  Each SpMV with [P | A'] is followed by one ncclAllReduce(sum) of n doubles
 * over NVLink before its fused epilogue runs; Ruiz column norms, the constraint-type flag
 * and the m-side residual scalars use allreduce(max / sum) of <= n doubles.
 *   osqp_cuda_nccl_unique_id : rank 0 creates the 128-byte NCCL id; the caller ships it to
 *                              the other ranks (any transport)
 *   osqp_cuda_set_partition  : every rank, before osqp_setup; world <= 1 leaves the mode.
 *                              Engines created afterwards are partitioned.
 * All ranks must make the same sequence of API calls (the collectives are matched by
 * order).  Returns 0 on success. */
c_int osqp_cuda_nccl_unique_id(unsigned char id128[128]);
c_int osqp_cuda_set_partition(c_int rank, c_int world, const unsigned char id128[128]);
c_int osqp_cuda_get_partition(c_int *rank, c_int *world, c_int *allreduces);
/* A world of ONE with the partitioned code path switched on (collectives are identities): lets a
 * single-GPU box run that path -- partial sums stored by the SpMV, epilogue in its own kernel --
 * and compare it with the ordinary solve.  on = 0 leaves the mode.                                */
c_int osqp_cuda_set_partition_loopback(c_int on);

/* ---- Lockstep batches (osqp_solve_batch of include/osqp.h on small problems; no counterpart
 * in the reference).  One call = `count` ADMM iterations (reference src/osqp.c:354-370) on
 * EVERY engine of the batch in one kernel launch with a thread block per problem, followed,
 * if do_resid, by the residual / infeasibility scalars of each (reference src/auxil.c:240-512,
 * same values as osqp_cuda_engine_residuals) in outs[i]: two launches, one D2H copy and one
 * synchronisation for the whole batch.  Engines must satisfy osqp_cuda_engine_is_fused.     */
c_int osqp_cuda_engine_is_fused(const osqp_cuda_engine *e);
c_int osqp_cuda_batch_step(osqp_cuda_engine **engines, c_int n, c_int count, c_float alpha,
                           c_int store_delta, c_int do_resid, osqp_cuda_resid **outs);
/* osqp_cuda_engine_cold_start (if cold; reference src/auxil.c:155-159) + osqp_cuda_engine_begin_solve for
 * every engine of a batch in one launch                                                              */
c_int osqp_cuda_batch_begin_solve(osqp_cuda_engine **engines, c_int n, c_int cold);
/* osqp_cuda_engine_get_solution (reference src/auxil.c:524-562 + src/scaling.c:177-192) for every
 * engine of a batch: one wait per group of problems, not per problem                          */
c_int osqp_cuda_batch_get_solutions(osqp_cuda_engine **engines, c_int n, c_float **xs, c_float **ys);

/* Upload P (upper-tri CSC), A (CSC), q, l, u; build CSR(A), CSR([P | A']) and
 * the CSC-slot -> CSR-slot maps on the device.  Replaces the host copies of
 * reference src/osqp.c:100-155.                                               */
c_int osqp_cuda_engine_create(osqp_cuda_engine **ep, c_int n, c_int m,
                              const csc *P, const csc *A, const c_float *q,
                              const c_float *l, const c_float *u);
void  osqp_cuda_engine_free(osqp_cuda_engine *e);

/* K11: Ruiz equilibration + cost scaling, reference src/scaling.c:44-156.
 * iters == 0 resets to the un-scaled problem.  Host outputs (any may be NULL):
 * D, Dinv (n), E, Einv (m), c, and the scaled q (n), l, u (m).                */
c_int osqp_cuda_engine_scale(osqp_cuda_engine *e, c_int iters,
                             c_float *D, c_float *Dinv, c_float *E, c_float *Einv,
                             c_float *c, c_float *q_s, c_float *l_s, c_float *u_s);

/* K5: classify constraints and build rho_vec (reference src/auxil.c:76-98),
 * then (re)build the Jacobi preconditioner diag(P)+sigma+sum rho a^2 and run
 * the curvature probe.  Creates the engine's cuda_pcg instance on first call.
 * Returns 0, OSQP_LINSYS_SOLVER_INIT_ERROR or OSQP_NONCVX_ERROR.              */
c_int osqp_cuda_engine_init_linsys(osqp_cuda_engine *e, c_float sigma, c_float rho,
                                   cuda_pcg_solver **solver_out);
/* reference src/osqp.c:1281-1332 (rho already clamped by the caller)          */
c_int osqp_cuda_engine_update_rho(osqp_cuda_engine *e, c_float rho);
/* reference src/auxil.c:100-142; *changed = 1 if any constraint type changed  */
c_int osqp_cuda_engine_retype_constraints(osqp_cuda_engine *e, c_float rho, c_int *changed);

/* K6 + K4 + K7: enqueue `count` ADMM iterations (reference src/osqp.c:354-370,
 * src/auxil.c:161-225).  No host synchronisation.  store_delta != 0 keeps
 * delta_x / delta_y of the LAST of them for the infeasibility tests.          */
c_int osqp_cuda_engine_iterate(osqp_cuda_engine *e, c_int count, c_float alpha,
                               c_int store_delta);
/* K9 + K10: one pass, synchronises, fills *out.  polished != 0 evaluates the
 * polished point instead of the iterate.                                      */
c_int osqp_cuda_engine_residuals(osqp_cuda_engine *e, osqp_cuda_resid *out, c_int polished);

/* reference src/auxil.c:155-159 */
c_int osqp_cuda_engine_cold_start(osqp_cuda_engine *e);
/* start of every osqp_solve: reset the CG launch policy (reproducible solves) */
c_int osqp_cuda_engine_begin_solve(osqp_cuda_engine *e);
/* reference src/osqp.c:942-1007: x <- Dinv x, z <- A x, y <- c Einv y.
 * x or y may be NULL (warm_start_x / warm_start_y).                           */
c_int osqp_cuda_engine_warm_start(osqp_cuda_engine *e, const c_float *x, const c_float *y);
/* reference src/osqp.c:765-940: raw host vectors in, scaled copies on device;
 * l or u may be NULL.  Scaled host copies out (may be NULL).                  */
c_int osqp_cuda_engine_update_q(osqp_cuda_engine *e, const c_float *q_new, c_float *q_s);
c_int osqp_cuda_engine_update_bounds(osqp_cuda_engine *e, const c_float *l_new,
                                     const c_float *u_new, c_float *l_s, c_float *u_s);
/* reference src/osqp.c:1012-1279, in three calls: (1) overwrite the raw values
 * (all, or by CSC index list; Px_new / Ax_new may be NULL = unchanged),
 * (2) osqp_cuda_engine_scale again, (3) osqp_cuda_engine_refactor: new Jacobi
 * diagonal + curvature probe, 0 or -1 (new K not positive definite).          */
c_int osqp_cuda_engine_update_matrices(osqp_cuda_engine *e,
                                       const c_float *Px_new, const c_int *Px_idx, c_int P_n,
                                       const c_float *Ax_new, const c_int *Ax_idx, c_int A_n);
c_int osqp_cuda_engine_refactor(osqp_cuda_engine *e);

/* reference src/auxil.c:524-562 + src/scaling.c:177-192: un-scaled x (n), y (m) */
c_int osqp_cuda_engine_get_solution(osqp_cuda_engine *e, c_float *x, c_float *y);
/* Allocation-free variant of osqp_cuda_engine_get_solution.  NO counterpart in the reference:
 * added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same outputs (HOST
 * x: n, y: m, either may be NULL); the un-scaled vectors are staged in a file-scope static
 * device array of OSQP_CUDA_NOALLOC_MAX doubles (2 MB, resident on every GPU that loads the
 * library until the process exits) instead of the engine's cudaMalloc'ed vectors.  The call
 * makes no cudaMalloc / cudaFree / allocating library call.  Returns 0; 2 (and a message on
 * stderr) when n + m > OSQP_CUDA_NOALLOC_MAX; 1 on a CUDA error.  All calls share the array:
 * it keeps the last call's contents (every call overwrites the slots it reads back), and two
 * calls that overlap on the GPU -- different engines on different streams or threads -- race. */
#define OSQP_CUDA_NOALLOC_MAX 262144
c_int osqp_cuda_engine_get_solution_noalloc(osqp_cuda_engine *e, c_float *x, c_float *y);
c_int osqp_cuda_noalloc_max(void);
/* scaled iterates; any may be NULL                                            */
c_int osqp_cuda_engine_get_iterates(osqp_cuda_engine *e, c_float *x, c_float *z, c_float *y);
/* certificates: delta_x (n, times D if unscale) / projected delta_y (m, times E) */
c_int osqp_cuda_engine_get_certificates(osqp_cuda_engine *e, c_float *delta_x,
                                        c_float *delta_y, c_int unscale);

/* K12: reference src/polish.c:212-350 on the device.  Outputs counts of
 * lower/upper-active rows; the polished point stays on the device until
 * accepted.  Returns 0 on success, nonzero if the reduced solve failed.       */
c_int osqp_cuda_engine_polish(osqp_cuda_engine *e, c_float delta, c_int refine_iter,
                              c_int *n_low, c_int *n_upp);
c_int osqp_cuda_engine_polish_accept(osqp_cuda_engine *e);

/* Kernel-level check point (tests): one SpMV launch of the solve path on caller data (HOST
 * pointers), through whichever kernel the solve path would take for this matrix.
 *   which 0: y (m) = A x (n)                          reference mat_vec, src/lin_alg.c:241-271
 *   which 1: y (n) = [Pf | A'] [x (n) ; x + n (m)]    reference mat_tpose_vec, src/lin_alg.c:273-322,
 *                                                     incl. the symmetric completion of P (src/auxil.c:296-306)
 *   which 2: y (2m) = [A x0 | A x1], x = [x0 ; x1]    two gathered vectors per pass (residual kernels)
 * The matrices are the engine's current device copies (scaled if scaling is on).            */
c_int osqp_cuda_engine_spmv(osqp_cuda_engine *e, c_int which, const c_float *x, c_float *y);

/* Counters and per-kernel timing (CUDA events on the engine's own stream).    */
c_int osqp_cuda_engine_get_stats(const osqp_cuda_engine *e, OSQPB200Stats *out);
c_int osqp_cuda_engine_profile(osqp_cuda_engine *e, c_int enable);
/* kernel ids: see OSQP_CUDA_K_* below; returns launches and total milliseconds */
c_int osqp_cuda_engine_profile_get(osqp_cuda_engine *e, c_int kernel_id,
                                   c_int *launches, c_float *total_ms, c_float *bytes_per_launch);
c_int osqp_cuda_engine_sync(osqp_cuda_engine *e);
/* Device-side stopwatch: CUDA events recorded on the engine's own stream (the one
 * every kernel of this engine is launched on).  stop synchronises the stream.   */
c_int osqp_cuda_engine_timer_start(osqp_cuda_engine *e);
c_int osqp_cuda_engine_timer_stop(osqp_cuda_engine *e, c_float *elapsed_ms);
/* PCG policy: the CG of one ADMM iteration stops at
 *   ||r||inf <= max(abs_floor, min(rel_tol * ||r0||inf, lambda * sqrt(pri_res * dua_res)))
 * (r0 = residual at the warm start x_k; pri/dua = scaled ADMM residuals of the last
 * check), or after max_iter iterations.                                       */
c_int osqp_cuda_engine_set_pcg_lambda(osqp_cuda_engine *e, c_float lambda);
c_int osqp_cuda_engine_set_pcg(osqp_cuda_engine *e, c_float rel_tol, c_float abs_floor,
                               c_int max_iter);

enum {
  OSQP_CUDA_K_SPMV_A      = 0,  /* w = rho .* (A p)          (inside PCG)       */
  OSQP_CUDA_K_SPMV_M      = 1,  /* Kp = [P|A'] [p;w] + sigma p (inside PCG)    */
  OSQP_CUDA_K_RHS         = 2,  /* r0 = -(q + [P|A'] [x;v]), z, p               */
  OSQP_CUDA_K_PCG_UPDATE  = 3,  /* x += a p, r -= a Kp, z = M^-1 r              */
  OSQP_CUDA_K_PCG_DIR     = 4,  /* p = z + b p                                  */
  OSQP_CUDA_K_ADMM_UPDATE = 5,  /* z~ = A x~ fused with x, z, y updates         */
  OSQP_CUDA_K_RESID_M     = 6,
  OSQP_CUDA_K_RESID_N     = 7,
  OSQP_CUDA_K_TILE_M      = 8,  /* column-blocked [P|A'] product alone (first kernel of 1 and 2) */
  OSQP_CUDA_K_TILE_A      = 9,  /* column-blocked A product alone (first kernel of 0; PARTIAL mode only) */
  OSQP_CUDA_K_ALLREDUCE   = 10, /* row-partitioned mode: ncclAllReduce of the [P|A'] partial sums   */
  OSQP_CUDA_K_APPLY_EPI   = 11, /* row-partitioned mode: fused row epilogue over the reduced sums   */
  OSQP_CUDA_K_COUNT       = 12
};

#ifdef __cplusplus
}
#endif
#endif /* OSQP_CUDA_PCG_H */
