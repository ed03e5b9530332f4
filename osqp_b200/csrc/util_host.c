/* util_host.c -- host support code of libosqp_b200.so: timers (reference
 * src/util.c:336-356), error reporting (src/error.c), input validation
 * (src/auxil.c:791-1065), status strings (src/auxil.c:646-679), progress
 * printing (src/util.c:42-236) and the SIGINT listener (src/ctrlc.c:54-80).   */
#include <signal.h>
#include <pthread.h>
#include "osqp_host.h"

/* ------------------------------------------------------------------ timers */
void osqp_tic(OSQPTimer *t) { clock_gettime(CLOCK_MONOTONIC, &t->tic); }

c_float osqp_toc(OSQPTimer *t) {
  clock_gettime(CLOCK_MONOTONIC, &t->toc);
  return (c_float)(t->toc.tv_sec - t->tic.tv_sec) +
         (c_float)(t->toc.tv_nsec - t->tic.tv_nsec) * 1e-9;
}

/* ------------------------------------------------------------------ errors */
const char *OSQP_ERROR_MESSAGE[] = {
  "Problem data validation.",
  "Solver settings validation.",
  "Linear system solver not available.\nTried to obtain it from shared library.",
  "Linear system solver initialization.",
  "KKT matrix factorization.\nThe problem seems to be non-convex.",
  "Memory allocation.",
  "Solver workspace not initialized.",
};

c_int osqp_error_(enum osqp_error_type code, const char *function_name) {
  c_print("ERROR in %s: %s\n", function_name, OSQP_ERROR_MESSAGE[code - 1]);
  return (c_int)code;
}

/* -------------------------------------------------------------- validation */
#define FAIL(...) do { c_eprint(__VA_ARGS__); return 1; } while (0)

c_int validate_data(const OSQPData *data) {
  c_int j, k;
  if (!data) FAIL("Missing data");
  if (!data->P) FAIL("Missing matrix P");
  if (!data->A) FAIL("Missing matrix A");
  if (!data->q) FAIL("Missing vector q");
  if (data->n <= 0 || data->m < 0)
    FAIL("n must be positive and m nonnegative; n = %i, m = %i", (int)data->n, (int)data->m);
  if (data->P->m != data->n) FAIL("P does not have dimension n x n with n = %i", (int)data->n);
  if (data->P->m != data->P->n) FAIL("P is not square");
  for (j = 0; j < data->n; j++)
    for (k = data->P->p[j]; k < data->P->p[j + 1]; k++)
      if (data->P->i[k] > j) FAIL("P is not upper triangular");
  if (data->A->m != data->m || data->A->n != data->n)
    FAIL("A does not have dimension %i x %i", (int)data->m, (int)data->n);
  for (j = 0; j < data->m; j++)
    if (data->l[j] > data->u[j])
      FAIL("Lower bound at index %d is greater than upper bound: %.4e > %.4e", (int)j,
           data->l[j], data->u[j]);
  return 0;
}

static c_int validate_linsys_solver(c_int linsys_solver) {
  /* the only backend in this library; QDLDL / Pardiso values are rejected, not emulated */
  return linsys_solver != CUDA_PCG_SOLVER;
}

c_int validate_settings(const OSQPSettings *s) {
  if (!s) FAIL("Missing settings!");
  if (s->scaling < 0) FAIL("scaling must be nonnegative");
  if (s->adaptive_rho != 0 && s->adaptive_rho != 1) FAIL("adaptive_rho must be either 0 or 1");
  if (s->adaptive_rho_interval < 0) FAIL("adaptive_rho_interval must be nonnegative");
  if (s->adaptive_rho_fraction <= 0) FAIL("adaptive_rho_fraction must be positive");
  if (s->adaptive_rho_tolerance < 1.0) FAIL("adaptive_rho_tolerance must be >= 1");
  if (s->polish_refine_iter < 0) FAIL("polish_refine_iter must be nonnegative");
  if (s->rho <= 0.0) FAIL("rho must be positive");
  if (s->sigma <= 0.0) FAIL("sigma must be positive");
  if (s->delta <= 0.0) FAIL("delta must be positive");
  if (s->max_iter <= 0) FAIL("max_iter must be positive");
  if (s->eps_abs < 0.0) FAIL("eps_abs must be nonnegative");
  if (s->eps_rel < 0.0) FAIL("eps_rel must be nonnegative");
  if (s->eps_rel == 0.0 && s->eps_abs == 0.0)
    FAIL("at least one of eps_abs and eps_rel must be positive");
  if (s->eps_prim_inf <= 0.0) FAIL("eps_prim_inf must be positive");
  if (s->eps_dual_inf <= 0.0) FAIL("eps_dual_inf must be positive");
  if (s->alpha <= 0.0 || s->alpha >= 2.0) FAIL("alpha must be strictly between 0 and 2");
  if (validate_linsys_solver(s->linsys_solver))
    FAIL("linsys_solver not recognized (this library provides CUDA_PCG_SOLVER = %d only)",
         (int)CUDA_PCG_SOLVER);
  if (s->verbose != 0 && s->verbose != 1) FAIL("verbose must be either 0 or 1");
  if (s->scaled_termination != 0 && s->scaled_termination != 1)
    FAIL("scaled_termination must be either 0 or 1");
  if (s->check_termination < 0) FAIL("check_termination must be nonnegative");
  if (s->warm_start != 0 && s->warm_start != 1) FAIL("warm_start must be either 0 or 1");
  if (s->time_limit < 0.0) FAIL("time_limit must be nonnegative");
  return 0;
}

/* ------------------------------------------------------------------ status */
void update_status(OSQPInfo *info, c_int status_val) {
  const char *txt = OSQP_NULL;
  info->status_val = status_val;
  switch (status_val) {
  case OSQP_SOLVED:                       txt = "solved"; break;
  case OSQP_SOLVED_INACCURATE:            txt = "solved inaccurate"; break;
  case OSQP_PRIMAL_INFEASIBLE:            txt = "primal infeasible"; break;
  case OSQP_PRIMAL_INFEASIBLE_INACCURATE: txt = "primal infeasible inaccurate"; break;
  case OSQP_UNSOLVED:                     txt = "unsolved"; break;
  case OSQP_DUAL_INFEASIBLE:              txt = "dual infeasible"; break;
  case OSQP_DUAL_INFEASIBLE_INACCURATE:   txt = "dual infeasible inaccurate"; break;
  case OSQP_MAX_ITER_REACHED:             txt = "maximum iterations reached"; break;
  case OSQP_TIME_LIMIT_REACHED:           txt = "run time limit reached"; break;
  case OSQP_SIGINT:                       txt = "interrupted"; break;
  case OSQP_NON_CVX:                      txt = "problem non convex"; break;
  default: break;
  }
  if (txt) { strncpy(info->status, txt, sizeof(info->status) - 1); info->status[31] = '\0'; }
}

void reset_info(OSQPInfo *info) {
  info->solve_time = 0.0;
  info->polish_time = 0.0;
  update_status(info, OSQP_UNSOLVED);
  info->rho_updates = 0;
}

c_int has_solution(const OSQPInfo *info) {
  return info->status_val != OSQP_PRIMAL_INFEASIBLE &&
         info->status_val != OSQP_PRIMAL_INFEASIBLE_INACCURATE &&
         info->status_val != OSQP_DUAL_INFEASIBLE &&
         info->status_val != OSQP_DUAL_INFEASIBLE_INACCURATE &&
         info->status_val != OSQP_NON_CVX;
}

/* ---------------------------------------------------------------- printing */
static void print_line(void) {
  c_print("-----------------------------------------------------------------\n");
}

void print_setup_header(const OSQPWorkspace *work) {
  const OSQPData *d = work->data;
  const OSQPSettings *s = work->settings;
  c_int nnz = d->P->p[d->P->n] + d->A->p[d->A->n];
  print_line();
  c_print("        OSQP-B200 v%s  -  ADMM QP solver, device-resident on B200\n", OSQP_VERSION);
  print_line();
  c_print("problem:  variables n = %i, constraints m = %i\n", (int)d->n, (int)d->m);
  c_print("          nnz(P) + nnz(A) = %i\n", (int)nnz);
  c_print("settings: linear system solver = %s", LINSYS_SOLVER_NAME[s->linsys_solver]);
  if (work->linsys_solver && work->linsys_solver->nthreads != 1)
    c_print(" (%d threads)", (int)work->linsys_solver->nthreads);
  c_print(",\n          eps_abs = %.1e, eps_rel = %.1e,\n", s->eps_abs, s->eps_rel);
  c_print("          eps_prim_inf = %.1e, eps_dual_inf = %.1e,\n", s->eps_prim_inf, s->eps_dual_inf);
  c_print("          rho = %.2e ", s->rho);
  if (s->adaptive_rho) c_print("(adaptive)");
  c_print(",\n          sigma = %.2e, alpha = %.2f, max_iter = %i\n", s->sigma, s->alpha, (int)s->max_iter);
  if (s->check_termination) c_print("          check_termination: on (interval %i),\n", (int)s->check_termination);
  else c_print("          check_termination: off,\n");
  if (s->time_limit) c_print("          time_limit: %.2e sec,\n", s->time_limit);
  if (s->scaling) c_print("          scaling: on, ");
  else c_print("          scaling: off, ");
  c_print("scaled_termination: %s\n", s->scaled_termination ? "on" : "off");
  c_print("          warm start: %s, polish: %s, time_limit: %s\n", s->warm_start ? "on" : "off",
          s->polish ? "on" : "off", s->time_limit ? "on" : "off");
  c_print("\n");
}

void print_header(void) {
  c_print("iter   objective    pri res    dua res    rho        time\n");
}

void print_summary(OSQPWorkspace *work) {
  const OSQPInfo *info = work->info;
  c_print("%4i  %12.4e  %9.2e  %9.2e  %9.2e  %9.2es\n", (int)info->iter, info->obj_val,
          info->pri_res, info->dua_res, work->settings->rho,
          work->first_run ? info->setup_time + info->solve_time : info->update_time + info->solve_time);
  work->summary_printed = 1;
}

void print_polish(OSQPWorkspace *work) {
  const OSQPInfo *info = work->info;
  c_print("plsh  %12.4e  %9.2e  %9.2e   --------  %9.2es\n", info->obj_val, info->pri_res,
          info->dua_res,
          (work->first_run ? info->setup_time : info->update_time) + info->solve_time + info->polish_time);
}

void print_footer(const OSQPInfo *info, c_int polish) {
  c_print("\nstatus:               %s\n", info->status);
  if (polish && info->status_val == OSQP_SOLVED) {
    if (info->status_polish == 1) c_print("solution polish:      successful\n");
    else if (info->status_polish < 0) c_print("solution polish:      unsuccessful\n");
  }
  c_print("number of iterations: %i\n", (int)info->iter);
  if (info->status_val == OSQP_SOLVED || info->status_val == OSQP_SOLVED_INACCURATE)
    c_print("optimal objective:    %.4f\n", info->obj_val);
  c_print("run time:             %.2es\n", info->run_time);
  c_print("optimal rho estimate: %.2e\n\n", info->rho_estimate);
}

/* ------------------------------------------------------------------ ctrl-c */
static volatile sig_atomic_t int_detected = 0;
static struct sigaction oact;
static void handle_ctrlc(int dummy) { (void)dummy; int_detected = 1; }

/* Ref-counted: osqp_solve_batch runs several solves concurrently, and the saved handler
 * must be the application's, not another solve's (reference src/ctrlc.c:54-80 is per call). */
static pthread_mutex_t int_lock = PTHREAD_MUTEX_INITIALIZER;
static int int_users = 0;
void osqp_start_interrupt_listener(void) {
  pthread_mutex_lock(&int_lock);
  if (int_users++ == 0) {
    struct sigaction act;
    int_detected = 0;
    act.sa_flags = 0;
    sigemptyset(&act.sa_mask);
    act.sa_handler = handle_ctrlc;
    sigaction(SIGINT, &act, &oact);
  }
  pthread_mutex_unlock(&int_lock);
}
void osqp_end_interrupt_listener(void) {
  pthread_mutex_lock(&int_lock);
  if (int_users > 0 && --int_users == 0) {
    struct sigaction act;
    sigaction(SIGINT, &oact, &act);
  }
  pthread_mutex_unlock(&int_lock);
}
int osqp_is_interrupted(void) { return int_detected; }
