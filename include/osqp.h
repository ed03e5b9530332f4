/*
 * osqp.h -- public C API of osqp-b200.
 *
 * One header instead of the reference's osqp.h / types.h / constants.h /
 * glob_opts.h quartet.  Every struct below is LAYOUT-COMPATIBLE with the
 * reference build configured as its CMake defaults configure it
 * (PRINTING, PROFILING, CTRLC, DLONG on; DFLOAT, EMBEDDED off), so a program
 * written against the reference re-links against libosqp_b200.so unchanged:
 *
 *   csc                    <- reference include/types.h:21-29
 *   OSQPScaling            <- include/types.h:44-51
 *   OSQPSolution           <- include/types.h:56-59
 *   OSQPInfo               <- include/types.h:65-91
 *   OSQPPolish             <- include/types.h:99-114
 *   OSQPData               <- include/types.h:125-133
 *   OSQPSettings           <- include/types.h:139-176
 *   OSQPWorkspace          <- include/types.h:182-288 (+ one trailing private ptr)
 *   struct linsys_solver   <- include/types.h:298-319
 *   status / error codes   <- include/constants.h:17-50
 *   defaults               <- include/constants.h:58-121
 *
 * What differs from the reference:
 *   - enum linsys_solver_type gains CUDA_PCG_SOLVER = 2, the only backend this
 *     library ships (no CPU fallback); osqp_set_default_settings() selects it.
 *   - iterates live in B200 HBM.  Host mirrors kept: data->{n,m,q,l,u},
 *     settings, scaling, info, solution, delta_x/delta_y (certificates).
 *     Workspace members that are pure device state in this build are NULL.
 */
#ifndef OSQP_B200_OSQP_H
#define OSQP_B200_OSQP_H

#ifdef __cplusplus
extern "C" {
#endif

/* ----------------------------------------------------------------- scalars */
typedef long long c_int;   /* reference glob_opts.h:79-83 with DLONG          */
typedef double    c_float; /* reference glob_opts.h:86-90 without DFLOAT      */

#define OSQP_VERSION "0.6.2-b200"

/* ------------------------------------------------------------ status codes */
#define OSQP_DUAL_INFEASIBLE_INACCURATE   (4)
#define OSQP_PRIMAL_INFEASIBLE_INACCURATE (3)
#define OSQP_SOLVED_INACCURATE            (2)
#define OSQP_SOLVED                       (1)
#define OSQP_MAX_ITER_REACHED             (-2)
#define OSQP_PRIMAL_INFEASIBLE            (-3)
#define OSQP_DUAL_INFEASIBLE              (-4)
#define OSQP_SIGINT                       (-5)
#define OSQP_TIME_LIMIT_REACHED           (-6)
#define OSQP_NON_CVX                      (-7)
#define OSQP_UNSOLVED                     (-10)

/* --------------------------------------------------- linear system solvers */
enum linsys_solver_type {
  QDLDL_SOLVER       = 0,  /* not built into this library                     */
  MKL_PARDISO_SOLVER = 1,  /* not built into this library                     */
  CUDA_PCG_SOLVER    = 2,  /* device-resident preconditioned CG (this repo)   */
  UNKNOWN_SOLVER     = 99
};
extern const char *LINSYS_SOLVER_NAME[];

/* -------------------------------------------------------------- error codes */
enum osqp_error_type {
  OSQP_DATA_VALIDATION_ERROR = 1,
  OSQP_SETTINGS_VALIDATION_ERROR,
  OSQP_LINSYS_SOLVER_LOAD_ERROR,
  OSQP_LINSYS_SOLVER_INIT_ERROR,
  OSQP_NONCVX_ERROR,
  OSQP_MEM_ALLOC_ERROR,
  OSQP_WORKSPACE_NOT_INIT_ERROR
};
extern const char *OSQP_ERROR_MESSAGE[];

/* ----------------------------------------------------------------- defaults */
#define RHO                   (0.1)
#define SIGMA                 (1E-06)
#define MAX_ITER              (4000)
#define EPS_ABS               (1E-3)
#define EPS_REL               (1E-3)
#define EPS_PRIM_INF          (1E-4)
#define EPS_DUAL_INF          (1E-4)
#define ALPHA                 (1.6)
#define LINSYS_SOLVER         (CUDA_PCG_SOLVER)
#define RHO_MIN               (1e-06)
#define RHO_MAX               (1e06)
#define RHO_EQ_OVER_RHO_INEQ  (1e03)
#define RHO_TOL               (1e-04)
#define DELTA                 (1E-6)
#define POLISH                (0)
#define POLISH_REFINE_ITER    (3)
#define VERBOSE               (1)
#define SCALED_TERMINATION    (0)
#define CHECK_TERMINATION     (25)
#define WARM_START            (1)
#define SCALING               (10)
#define MIN_SCALING           (1e-04)
#define MAX_SCALING           (1e+04)
#define ADAPTIVE_RHO                      (1)
#define ADAPTIVE_RHO_INTERVAL             (0)
#define ADAPTIVE_RHO_FRACTION             (0.4)
#define ADAPTIVE_RHO_MULTIPLE_TERMINATION (4)
#define ADAPTIVE_RHO_FIXED                (100)
#define ADAPTIVE_RHO_TOLERANCE            (5)
#define TIME_LIMIT            (0)
#define PRINT_INTERVAL        200

#ifndef OSQP_NULL
# define OSQP_NULL 0
#endif
/* The reference's "NaN" is the NUMBER 2143289344.0 (constants.h:95) and its
 * tests compare against it with ==; kept bit-for-bit.                        */
#ifndef OSQP_NAN
# define OSQP_NAN ((c_float)0x7fc00000UL)
#endif
#ifndef OSQP_INFTY
# define OSQP_INFTY ((c_float)1e30)
#endif
#ifndef OSQP_DIVISION_TOL
# define OSQP_DIVISION_TOL ((c_float)1.0 / OSQP_INFTY)
#endif

/* -------------------------------------------------------------------- types */
typedef struct {
  c_int    nzmax; /* allocated entries                                        */
  c_int    m;     /* rows                                                     */
  c_int    n;     /* columns                                                  */
  c_int   *p;     /* column pointers, n+1                                     */
  c_int   *i;     /* row indices, nzmax                                       */
  c_float *x;     /* values, nzmax                                            */
  c_int    nz;    /* -1 for compressed-column                                 */
} csc;

typedef struct linsys_solver LinSysSolver;
typedef struct OSQP_TIMER    OSQPTimer;

typedef struct {
  c_float  c;
  c_float *D;
  c_float *E;
  c_float  cinv;
  c_float *Dinv;
  c_float *Einv;
} OSQPScaling;

typedef struct {
  c_float *x;
  c_float *y;
} OSQPSolution;

typedef struct {
  c_int   iter;
  char    status[32];
  c_int   status_val;
  c_int   status_polish;
  c_float obj_val;
  c_float pri_res;
  c_float dua_res;
  c_float setup_time;
  c_float solve_time;
  c_float update_time;
  c_float polish_time;
  c_float run_time;
  c_int   rho_updates;
  c_float rho_estimate;
} OSQPInfo;

typedef struct {
  csc     *Ared;      /* never materialised here: the device polish masks rows */
  c_int    n_low;
  c_int    n_upp;
  c_int   *A_to_Alow;
  c_int   *A_to_Aupp;
  c_int   *Alow_to_A;
  c_int   *Aupp_to_A;
  c_float *x;
  c_float *z;
  c_float *y;
  c_float  obj_val;
  c_float  pri_res;
  c_float  dua_res;
} OSQPPolish;

typedef struct {
  c_int    n;
  c_int    m;
  csc     *P;
  csc     *A;
  c_float *q;
  c_float *l;
  c_float *u;
} OSQPData;

typedef struct {
  c_float rho;
  c_float sigma;
  c_int   scaling;
  c_int   adaptive_rho;
  c_int   adaptive_rho_interval;
  c_float adaptive_rho_tolerance;
  c_float adaptive_rho_fraction;
  c_int   max_iter;
  c_float eps_abs;
  c_float eps_rel;
  c_float eps_prim_inf;
  c_float eps_dual_inf;
  c_float alpha;
  enum linsys_solver_type linsys_solver;
  c_float delta;
  c_int   polish;
  c_int   polish_refine_iter;
  c_int   verbose;
  c_int   scaled_termination;
  c_int   check_termination;
  c_int   warm_start;
  c_float time_limit;
} OSQPSettings;

typedef struct {
  OSQPData     *data;          /* host mirror: n, m, q, l, u (scaled); P/A: p only */
  LinSysSolver *linsys_solver; /* the cuda_pcg instance                       */
  OSQPPolish   *pol;
  c_float *rho_vec;            /* device state here: NULL                     */
  c_float *rho_inv_vec;        /* NULL                                        */
  c_int   *constr_type;        /* NULL                                        */
  c_float *x;                  /* NULL (device)                               */
  c_float *y;                  /* NULL (device)                               */
  c_float *z;                  /* NULL (device)                               */
  c_float *xz_tilde;           /* NULL                                        */
  c_float *x_prev;             /* NULL                                        */
  c_float *z_prev;             /* NULL                                        */
  c_float *Ax;                 /* NULL                                        */
  c_float *Px;                 /* NULL                                        */
  c_float *Aty;                /* NULL                                        */
  c_float *delta_y;            /* host: primal infeasibility certificate      */
  c_float *Atdelta_y;          /* NULL                                        */
  c_float *delta_x;            /* host: dual infeasibility certificate        */
  c_float *Pdelta_x;           /* NULL                                        */
  c_float *Adelta_x;           /* NULL                                        */
  c_float *D_temp;             /* NULL                                        */
  c_float *D_temp_A;           /* NULL                                        */
  c_float *E_temp;             /* NULL                                        */
  OSQPSettings *settings;
  OSQPScaling  *scaling;
  OSQPSolution *solution;
  OSQPInfo     *info;
  OSQPTimer    *timer;
  c_int first_run;
  c_int clear_update_time;
  c_int rho_update_from_solve;
  c_int summary_printed;
  void *b200;                  /* private: device engine + bookkeeping        */
} OSQPWorkspace;

struct linsys_solver {
  enum linsys_solver_type type;
  c_int (*solve)(LinSysSolver *self, c_float *b);
  void  (*free)(LinSysSolver *self);
  c_int (*update_matrices)(LinSysSolver *s, const csc *P, const csc *A);
  c_int (*update_rho_vec)(LinSysSolver *s, const c_float *rho_vec);
  c_int nthreads;
};

/* ---------------------------------------------------------------- main API
 * Same names, argument meaning and return codes as reference include/osqp.h. */

/* reference src/osqp.c:24-71 (linsys_solver default is CUDA_PCG_SOLVER here) */
void  osqp_set_default_settings(OSQPSettings *settings);

/* reference src/osqp.c:76-283.  0 on success, else an osqp_error_type.
 * Uploads P, A, q, l, u to HBM, equilibrates on the device, builds CSR(A),
 * CSR([P | A']) and the Jacobi preconditioner.  No factorisation.            */
c_int osqp_setup(OSQPWorkspace **workp, const OSQPData *data,
                 const OSQPSettings *settings);

/* reference src/osqp.c:288-654.  Device-resident ADMM loop.                  */
c_int osqp_solve(OSQPWorkspace *work);

/* reference src/osqp.c:659-757 */
c_int osqp_cleanup(OSQPWorkspace *work);

/* ----------------------------------------------------------- data updates */
c_int osqp_update_lin_cost(OSQPWorkspace *work, const c_float *q_new);        /* osqp.c:765 */
c_int osqp_update_bounds(OSQPWorkspace *work, const c_float *l_new,
                         const c_float *u_new);                               /* osqp.c:797 */
c_int osqp_update_lower_bound(OSQPWorkspace *work, const c_float *l_new);     /* osqp.c:850 */
c_int osqp_update_upper_bound(OSQPWorkspace *work, const c_float *u_new);     /* osqp.c:896 */
c_int osqp_warm_start(OSQPWorkspace *work, const c_float *x, const c_float *y); /* osqp.c:942 */
c_int osqp_warm_start_x(OSQPWorkspace *work, const c_float *x);               /* osqp.c:969 */
c_int osqp_warm_start_y(OSQPWorkspace *work, const c_float *y);               /* osqp.c:990 */
c_int osqp_update_P(OSQPWorkspace *work, const c_float *Px_new,
                    const c_int *Px_new_idx, c_int P_new_n);                  /* osqp.c:1012 */
c_int osqp_update_A(OSQPWorkspace *work, const c_float *Ax_new,
                    const c_int *Ax_new_idx, c_int A_new_n);                  /* osqp.c:1093 */
c_int osqp_update_P_A(OSQPWorkspace *work, const c_float *Px_new,
                      const c_int *Px_new_idx, c_int P_new_n,
                      const c_float *Ax_new, const c_int *Ax_new_idx,
                      c_int A_new_n);                                         /* osqp.c:1173 */
c_int osqp_update_rho(OSQPWorkspace *work, c_float rho_new);                  /* osqp.c:1281 */

/* ------------------------------------------------------- settings updates */
c_int osqp_update_max_iter(OSQPWorkspace *work, c_int max_iter_new);
c_int osqp_update_eps_abs(OSQPWorkspace *work, c_float eps_abs_new);
c_int osqp_update_eps_rel(OSQPWorkspace *work, c_float eps_rel_new);
c_int osqp_update_eps_prim_inf(OSQPWorkspace *work, c_float eps_prim_inf_new);
c_int osqp_update_eps_dual_inf(OSQPWorkspace *work, c_float eps_dual_inf_new);
c_int osqp_update_alpha(OSQPWorkspace *work, c_float alpha_new);
c_int osqp_update_warm_start(OSQPWorkspace *work, c_int warm_start_new);
c_int osqp_update_scaled_termination(OSQPWorkspace *work, c_int scaled_termination_new);
c_int osqp_update_check_termination(OSQPWorkspace *work, c_int check_termination_new);
c_int osqp_update_delta(OSQPWorkspace *work, c_float delta_new);
c_int osqp_update_polish(OSQPWorkspace *work, c_int polish_new);
c_int osqp_update_polish_refine_iter(OSQPWorkspace *work, c_int polish_refine_iter_new);
c_int osqp_update_verbose(OSQPWorkspace *work, c_int verbose_new);
c_int osqp_update_time_limit(OSQPWorkspace *work, c_float time_limit_new);

/* ------------------------------------------- linear-system plugin boundary
 * reference include/lin_sys.h:17-48, src/lin_sys.c:15-75.  For
 * CUDA_PCG_SOLVER load/unload dlopen()/dlclose() libosqp_cuda_pcg.so through
 * the shared-library handler (ref-counted), exactly as the reference does
 * for MKL Pardiso (lin_sys/direct/pardiso/pardiso_loader.c:63-102).          */
c_int load_linsys_solver(enum linsys_solver_type linsys_solver);
c_int unload_linsys_solver(enum linsys_solver_type linsys_solver);
c_int init_linsys_solver(LinSysSolver **s, const csc *P, const csc *A,
                         c_float sigma, const c_float *rho_vec,
                         enum linsys_solver_type linsys_solver, c_int polish);

/* ----------------------------------------------- helpers kept from cs.h/util.h
 * csc_matrix wraps caller arrays without copying (reference src/cs.c:12-26). */
csc  *csc_matrix(c_int m, c_int n, c_int nzmax, c_float *x, c_int *i, c_int *p);
void  cold_start(OSQPWorkspace *work);   /* reference src/auxil.c:155-159      */

/* -------------------------------------------------- additive (not in ref.)
 * Counters of the device path since setup: ADMM iterations, PCG iterations,
 * kernel launches.  Used by bench.py for N_pcg and gpu_launches.             */
typedef struct {
  c_int admm_iters;
  c_int pcg_iters;
  c_int kernel_launches;
  c_int rho_updates;
  c_int pcg_maxed;       /* solves that hit the PCG iteration cap             */
  c_int pcg_breakdowns;  /* solves stopped by p'Kp <= 0 (or NaN)              */
  c_int sell_launches;   /* SpMV launches that streamed the sliced-ELL mirror */
  c_int tile_launches;   /* SpMV launches that went through a column-blocked mirror */
  c_int pcg_noncvx;      /* 1: the CG met non-positive curvature (K not positive definite) */
  c_int slack_vars;      /* variables eliminated exactly from the CG's system (csrc/slack.cuh) */
} OSQPB200Stats;
c_int osqp_b200_get_stats(const OSQPWorkspace *work, OSQPB200Stats *out);
/* Scaled iterates (x: n, z: m, y: m) copied D2H on request; any may be NULL.  */
c_int osqp_b200_get_iterates(const OSQPWorkspace *work, c_float *x, c_float *z, c_float *y);
/* Opaque handle of the device engine (osqp_cuda_engine *) for the profiling
 * entry points of osqp_cuda_pcg.h.                                           */
void *osqp_b200_engine(const OSQPWorkspace *work);
/* Path of the CUDA plugin library to dlopen (default: next to this library). */
void  osqp_b200_set_plugin_path(const char *path);
/* Batches of independent QPs on one GPU (BASELINE config 4; the reference has no batch entry
 * point).  Each element is an ordinary workspace; up to `nthreads` (<= 256) of them are solved
 * concurrently, each on its own CUDA stream, with results identical to osqp_solve's.
 * Returns the number of elements whose solve failed. */
c_int osqp_solve_batch(OSQPWorkspace **works, c_int count, c_int nthreads);

#ifdef __cplusplus
}
#endif
#endif /* OSQP_B200_OSQP_H */
