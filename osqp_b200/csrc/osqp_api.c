/* osqp_api.c -- osqp_setup / osqp_solve / osqp_update_* / osqp_cleanup on top of the
 * device-resident engine.  Same entry points, argument meaning and return codes as
 * reference src/osqp.c; what runs on the host here is control flow only (validation,
 * termination decisions from a handful of scalars, adaptive-rho policy, timing,
 * printing).  All vector and matrix work is behind cuda_api.*                    */
#include "osqp_host.h"

#define PRIV(w) ((OSQPB200Priv *)((w)->b200))
#define ENGINE(w) (PRIV(w)->engine)

/* ---------------------------------------------------------------- defaults */
void osqp_set_default_settings(OSQPSettings *s) {   /* reference src/osqp.c:24-71 */
  s->rho = (c_float)RHO;
  s->sigma = (c_float)SIGMA;
  s->scaling = SCALING;
  s->adaptive_rho = ADAPTIVE_RHO;
  s->adaptive_rho_interval = ADAPTIVE_RHO_INTERVAL;
  s->adaptive_rho_tolerance = (c_float)ADAPTIVE_RHO_TOLERANCE;
  s->adaptive_rho_fraction = (c_float)ADAPTIVE_RHO_FRACTION;
  s->max_iter = MAX_ITER;
  s->eps_abs = (c_float)EPS_ABS;
  s->eps_rel = (c_float)EPS_REL;
  s->eps_prim_inf = (c_float)EPS_PRIM_INF;
  s->eps_dual_inf = (c_float)EPS_DUAL_INF;
  s->alpha = (c_float)ALPHA;
  s->linsys_solver = LINSYS_SOLVER;
  s->delta = DELTA;
  s->polish = POLISH;
  s->polish_refine_iter = POLISH_REFINE_ITER;
  s->verbose = VERBOSE;
  s->scaled_termination = SCALED_TERMINATION;
  s->check_termination = CHECK_TERMINATION;
  s->warm_start = WARM_START;
  s->time_limit = TIME_LIMIT;
}

/* ------------------------------------------------------------- csc helpers */
csc *csc_matrix(c_int m, c_int n, c_int nzmax, c_float *x, c_int *i, c_int *p) {
  csc *M = (csc *)malloc(sizeof(csc));   /* reference src/cs.c:12-26: wraps, no copy */
  if (!M) return OSQP_NULL;
  M->m = m; M->n = n; M->nz = -1; M->nzmax = nzmax; M->x = x; M->i = i; M->p = p;
  return M;
}
/* host mirror of a matrix: dimensions + column pointers; values/indices live in HBM */
static csc *csc_mirror(const csc *src) {
  csc *M = (csc *)calloc(1, sizeof(csc));
  if (!M) return OSQP_NULL;
  M->m = src->m; M->n = src->n; M->nz = -1; M->nzmax = src->p[src->n];
  M->p = (c_int *)malloc((size_t)(src->n + 1) * sizeof(c_int));
  if (!M->p) { free(M); return OSQP_NULL; }
  memcpy(M->p, src->p, (size_t)(src->n + 1) * sizeof(c_int));
  return M;
}
static void csc_mirror_free(csc *M) {
  if (!M) return;
  if (M->p) free(M->p);
  free(M);
}
static c_float *vec_dup(const c_float *a, c_int len) {
  c_float *b = (c_float *)malloc((size_t)(len > 0 ? len : 1) * sizeof(c_float));
  if (b && len > 0) memcpy(b, a, (size_t)len * sizeof(c_float));
  return b;
}

void cold_start(OSQPWorkspace *work) {   /* reference src/auxil.c:155-159 */
  if (work && work->b200 && ENGINE(work)) cuda_api.engine_cold_start(ENGINE(work));
}

/* ------------------------------------------------------------------- setup */
c_int osqp_setup(OSQPWorkspace **workp, const OSQPData *data, const OSQPSettings *settings) {
  c_int exitflag, n, m;
  OSQPWorkspace *work;
  OSQPB200Priv *priv;

  if (validate_data(data)) return osqp_error(OSQP_DATA_VALIDATION_ERROR);
  if (validate_settings(settings)) return osqp_error(OSQP_SETTINGS_VALIDATION_ERROR);

  work = (OSQPWorkspace *)calloc(1, sizeof(OSQPWorkspace));
  if (!work) return osqp_error(OSQP_MEM_ALLOC_ERROR);
  *workp = work;
  n = data->n; m = data->m;

  work->timer = (OSQPTimer *)malloc(sizeof(OSQPTimer));
  if (!work->timer) return osqp_error(OSQP_MEM_ALLOC_ERROR);
  osqp_tic(work->timer);

  priv = (OSQPB200Priv *)calloc(1, sizeof(OSQPB200Priv));
  if (!priv) return osqp_error(OSQP_MEM_ALLOC_ERROR);
  work->b200 = priv;

  /* host mirrors of the problem data */
  work->data = (OSQPData *)calloc(1, sizeof(OSQPData));
  if (!work->data) return osqp_error(OSQP_MEM_ALLOC_ERROR);
  work->data->n = n; work->data->m = m;
  work->data->P = csc_mirror(data->P);
  work->data->A = csc_mirror(data->A);
  work->data->q = vec_dup(data->q, n);
  work->data->l = vec_dup(data->l, m);
  work->data->u = vec_dup(data->u, m);
  if (!work->data->P || !work->data->A || !work->data->q || !work->data->l || !work->data->u)
    return osqp_error(OSQP_MEM_ALLOC_ERROR);

  /* certificates are the only iterate-sized vectors mirrored on the host */
  work->delta_x = (c_float *)calloc((size_t)n, sizeof(c_float));
  work->delta_y = (c_float *)calloc((size_t)(m > 0 ? m : 1), sizeof(c_float));
  if (!work->delta_x || !work->delta_y) return osqp_error(OSQP_MEM_ALLOC_ERROR);

  work->settings = (OSQPSettings *)malloc(sizeof(OSQPSettings));
  if (!work->settings) return osqp_error(OSQP_MEM_ALLOC_ERROR);
  memcpy(work->settings, settings, sizeof(OSQPSettings));

  /* load the CUDA backend (dlopen; fails loudly when no B200 is usable) */
  if (load_linsys_solver(work->settings->linsys_solver))
    return osqp_error(OSQP_LINSYS_SOLVER_LOAD_ERROR);
  priv->plugin_loaded = 1;

  if (getenv("OSQP_B200_TIMING")) c_print("[timing] host mirrors + plugin load: %.3f ms\n", 1e3 * osqp_toc(work->timer));
  /* upload + device format conversion */
  if (cuda_api.engine_create(&priv->engine, n, m, data->P, data->A, data->q, data->l, data->u))
    return osqp_error(OSQP_LINSYS_SOLVER_INIT_ERROR);

  if (getenv("OSQP_B200_TIMING")) { cuda_api.engine_sync(priv->engine); c_print("[timing] engine_create done: %.3f ms\n", 1e3 * osqp_toc(work->timer)); }
  /* scaling (K11); the host keeps D, E, c and the scaled q, l, u */
  if (settings->scaling) {
    work->scaling = (OSQPScaling *)calloc(1, sizeof(OSQPScaling));
    if (!work->scaling) return osqp_error(OSQP_MEM_ALLOC_ERROR);
    work->scaling->D = (c_float *)malloc((size_t)n * sizeof(c_float));
    work->scaling->Dinv = (c_float *)malloc((size_t)n * sizeof(c_float));
    work->scaling->E = (c_float *)malloc((size_t)(m > 0 ? m : 1) * sizeof(c_float));
    work->scaling->Einv = (c_float *)malloc((size_t)(m > 0 ? m : 1) * sizeof(c_float));
    if (!work->scaling->D || !work->scaling->Dinv || !work->scaling->E || !work->scaling->Einv)
      return osqp_error(OSQP_MEM_ALLOC_ERROR);
    if (cuda_api.engine_scale(priv->engine, settings->scaling, work->scaling->D,
                              work->scaling->Dinv, work->scaling->E, work->scaling->Einv,
                              &work->scaling->c, work->data->q, work->data->l, work->data->u))
      return osqp_error(OSQP_LINSYS_SOLVER_INIT_ERROR);
    work->scaling->cinv = 1. / work->scaling->c;
  } else {
    work->scaling = OSQP_NULL;
    if (cuda_api.engine_scale(priv->engine, 0, OSQP_NULL, OSQP_NULL, OSQP_NULL, OSQP_NULL,
                              OSQP_NULL, OSQP_NULL, OSQP_NULL, OSQP_NULL))
      return osqp_error(OSQP_LINSYS_SOLVER_INIT_ERROR);
  }

  if (getenv("OSQP_B200_TIMING")) { cuda_api.engine_sync(priv->engine); c_print("[timing] scaling done: %.3f ms\n", 1e3 * osqp_toc(work->timer)); }
  /* rho_vec, Jacobi preconditioner, curvature probe (replaces the factorisation) */
  work->settings->rho = c_min(c_max(work->settings->rho, RHO_MIN), RHO_MAX);
  exitflag = cuda_api.engine_init_linsys(priv->engine, work->settings->sigma, work->settings->rho,
                                         (cuda_pcg_solver **)&work->linsys_solver);
  if (exitflag) return osqp_error((enum osqp_error_type)exitflag);

  if (getenv("OSQP_B200_TIMING")) { cuda_api.engine_sync(priv->engine); c_print("[timing] init_linsys done: %.3f ms\n", 1e3 * osqp_toc(work->timer)); }
  work->pol = (OSQPPolish *)calloc(1, sizeof(OSQPPolish));
  work->solution = (OSQPSolution *)calloc(1, sizeof(OSQPSolution));
  work->info = (OSQPInfo *)calloc(1, sizeof(OSQPInfo));
  if (!work->pol || !work->solution || !work->info) return osqp_error(OSQP_MEM_ALLOC_ERROR);
  work->solution->x = (c_float *)calloc((size_t)n, sizeof(c_float));
  work->solution->y = (c_float *)calloc((size_t)(m > 0 ? m : 1), sizeof(c_float));
  if (!work->solution->x || !work->solution->y) return osqp_error(OSQP_MEM_ALLOC_ERROR);

  work->info->status_polish = 0;
  update_status(work->info, OSQP_UNSOLVED);
  work->info->solve_time = 0.0;
  work->info->update_time = 0.0;
  work->info->polish_time = 0.0;
  work->info->run_time = 0.0;
  cuda_api.engine_sync(priv->engine);
  work->info->setup_time = osqp_toc(work->timer);
  work->first_run = 1;
  work->clear_update_time = 0;
  work->rho_update_from_solve = 0;
  work->info->rho_updates = 0;
  work->info->rho_estimate = work->settings->rho;

  if (work->settings->verbose) print_setup_header(work);
  work->summary_printed = 0;
  return 0;
}

/* ----------------------------------------------------------------- cleanup */
c_int osqp_cleanup(OSQPWorkspace *work) {   /* reference src/osqp.c:659-757 */
  c_int exitflag = 0;
  if (!work) return 0;
  if (work->data) {
    csc_mirror_free(work->data->P);
    csc_mirror_free(work->data->A);
    if (work->data->q) free(work->data->q);
    if (work->data->l) free(work->data->l);
    if (work->data->u) free(work->data->u);
    free(work->data);
  }
  if (work->scaling) {
    if (work->scaling->D) free(work->scaling->D);
    if (work->scaling->Dinv) free(work->scaling->Dinv);
    if (work->scaling->E) free(work->scaling->E);
    if (work->scaling->Einv) free(work->scaling->Einv);
    free(work->scaling);
  }
  if (work->linsys_solver && work->linsys_solver->free)
    work->linsys_solver->free(work->linsys_solver);
  if (work->b200) {
    OSQPB200Priv *priv = PRIV(work);
    if (priv->engine) cuda_api.engine_free(priv->engine);
    if (priv->plugin_loaded && work->settings)
      exitflag = unload_linsys_solver(work->settings->linsys_solver);
    free(priv);
  }
  if (work->pol) free(work->pol);
  if (work->delta_x) free(work->delta_x);
  if (work->delta_y) free(work->delta_y);
  if (work->settings) free(work->settings);
  if (work->solution) {
    if (work->solution->x) free(work->solution->x);
    if (work->solution->y) free(work->solution->y);
    free(work->solution);
  }
  if (work->info) free(work->info);
  if (work->timer) free(work->timer);
  free(work);
  return exitflag;
}

/* ------------------------------------------------ residuals and termination */
/* reference src/auxil.c:564-629 (update_info): one device pass, ~30 scalars D2H */
/* info fields from the residual scalars already in PRIV(work)->res */
static void update_info_apply(OSQPWorkspace *work, c_int iter, c_int compute_objective, c_int polish) {
  OSQPB200Priv *priv = PRIV(work);
  osqp_cuda_resid *R = &priv->res;
  const c_int unsc = work->settings->scaling && !work->settings->scaled_termination;
  c_float obj, pri, dua;

  priv->res_valid = 1;
  obj = 0.5 * R->xPx + R->qx;
  if (work->settings->scaling) obj *= work->scaling->cinv;
  pri = (work->data->m == 0) ? 0. : (unsc ? R->pri_res_u : R->pri_res_s);
  dua = unsc ? work->scaling->cinv * R->dua_res_u : R->dua_res_s;
  if (polish) {
    work->pol->obj_val = obj;
    work->pol->pri_res = pri;
    work->pol->dua_res = dua;
    work->info->polish_time = osqp_toc(work->timer);
  } else {
    work->info->iter = iter;
    if (compute_objective) work->info->obj_val = obj;
    work->info->pri_res = pri;
    work->info->dua_res = dua;
    work->info->solve_time = osqp_toc(work->timer);
  }
  work->summary_printed = 0;
}

static c_int update_info(OSQPWorkspace *work, c_int iter, c_int compute_objective, c_int polish) {
  if (cuda_api.engine_residuals(PRIV(work)->engine, &PRIV(work)->res, polish)) return 1;
  update_info_apply(work, iter, compute_objective, polish);
  return 0;
}

static c_float compute_obj_val(OSQPWorkspace *work) {   /* reference src/auxil.c:227-238 */
  const osqp_cuda_resid *R = &PRIV(work)->res;
  c_float obj = 0.5 * R->xPx + R->qx;
  if (work->settings->scaling) obj *= work->scaling->cinv;
  return obj;
}

/* reference src/auxil.c:361-424, evaluated from the device scalars */
static c_int is_primal_infeasible(const OSQPWorkspace *work, c_float eps_prim_inf) {
  const osqp_cuda_resid *R = &PRIV(work)->res;
  const c_int unsc = work->settings->scaling && !work->settings->scaled_termination;
  const c_float norm_dy = unsc ? R->dy_norm_u : R->dy_norm_s;
  if (norm_dy > OSQP_DIVISION_TOL) {
    if (R->ineq_lhs < eps_prim_inf * norm_dy)
      return (unsc ? R->Atdy_norm_u : R->Atdy_norm_s) < eps_prim_inf * norm_dy;
  }
  return 0;
}

/* reference src/auxil.c:426-512; the row-wise early-exit loop is a max/min reduction */
static c_int is_dual_infeasible(const OSQPWorkspace *work, c_float eps_dual_inf) {
  const osqp_cuda_resid *R = &PRIV(work)->res;
  const c_int unsc = work->settings->scaling && !work->settings->scaled_termination;
  const c_float norm_dx = unsc ? R->dx_norm_u : R->dx_norm_s;
  const c_float cost_scaling = unsc ? work->scaling->c : 1.0;
  if (norm_dx > OSQP_DIVISION_TOL) {
    if (R->q_dx < cost_scaling * eps_dual_inf * norm_dx) {
      if ((unsc ? R->Pdx_norm_u : R->Pdx_norm_s) < cost_scaling * eps_dual_inf * norm_dx) {
        const c_float maxU = unsc ? R->Adx_maxU_u : R->Adx_maxU_s;
        const c_float minL = unsc ? R->Adx_minL_u : R->Adx_minL_s;
        if (maxU > eps_dual_inf * norm_dx) return 0;
        if (minL < -eps_dual_inf * norm_dx) return 0;
        return 1;
      }
    }
  }
  return 0;
}

/* reference src/auxil.c:681-786 */
static c_int check_termination(OSQPWorkspace *work, c_int approximate) {
  const osqp_cuda_resid *R = &PRIV(work)->res;
  const c_int unsc = work->settings->scaling && !work->settings->scaled_termination;
  c_float eps_prim, eps_dual, max_rel;
  c_float eps_abs = work->settings->eps_abs, eps_rel = work->settings->eps_rel;
  c_float eps_prim_inf = work->settings->eps_prim_inf, eps_dual_inf = work->settings->eps_dual_inf;
  c_int prim_res_check = 0, dual_res_check = 0, prim_inf_check = 0, dual_inf_check = 0;

  /* diverging (or NaN) residuals: the problem is most likely non-convex (reference
   * src/auxil.c:699-706).  The device CG sees it directly: negative curvature of K. */
  if (work->info->pri_res > OSQP_INFTY || work->info->dua_res > OSQP_INFTY ||
      work->info->pri_res != work->info->pri_res || work->info->dua_res != work->info->dua_res ||
      R->pcg_noncvx != 0.0) {
    update_status(work->info, OSQP_NON_CVX);
    work->info->obj_val = OSQP_NAN;
    return 1;
  }
  if (approximate) { eps_abs *= 10; eps_rel *= 10; eps_prim_inf *= 10; eps_dual_inf *= 10; }

  if (work->data->m == 0) {
    prim_res_check = 1;
  } else {
    max_rel = unsc ? c_max(R->z_norm_u, R->Ax_norm_u) : c_max(R->z_norm_s, R->Ax_norm_s);
    eps_prim = eps_abs + eps_rel * max_rel;
    if (work->info->pri_res < eps_prim) prim_res_check = 1;
    else prim_inf_check = is_primal_infeasible(work, eps_prim_inf);
  }
  if (unsc) {
    max_rel = c_max(c_max(R->q_norm_u, R->Aty_norm_u), R->Px_norm_u) * work->scaling->cinv;
  } else {
    max_rel = c_max(c_max(R->q_norm_s, R->Aty_norm_s), R->Px_norm_s);
  }
  eps_dual = eps_abs + eps_rel * max_rel;
  if (work->info->dua_res < eps_dual) dual_res_check = 1;
  else dual_inf_check = is_dual_infeasible(work, eps_dual_inf);

  if (prim_res_check && dual_res_check) {
    update_status(work->info, approximate ? OSQP_SOLVED_INACCURATE : OSQP_SOLVED);
    return 1;
  }
  if (prim_inf_check) {
    update_status(work->info, approximate ? OSQP_PRIMAL_INFEASIBLE_INACCURATE : OSQP_PRIMAL_INFEASIBLE);
    work->info->obj_val = OSQP_INFTY;
    return 1;
  }
  if (dual_inf_check) {
    update_status(work->info, approximate ? OSQP_DUAL_INFEASIBLE_INACCURATE : OSQP_DUAL_INFEASIBLE);
    work->info->obj_val = -OSQP_INFTY;
    return 1;
  }
  return 0;
}

/* reference src/auxil.c:13-52: uses the SCALED residuals and norms */
static c_float compute_rho_estimate(const OSQPWorkspace *work) {
  const osqp_cuda_resid *R = &PRIV(work)->res;
  c_float pri_res = R->pri_res_s, dua_res = R->dua_res_s, rho_estimate;
  c_float pri_norm = c_max(R->z_norm_s, R->Ax_norm_s);
  c_float dua_norm = c_max(c_max(R->q_norm_s, R->Aty_norm_s), R->Px_norm_s);
  if (work->data->m == 0) { pri_res = 0.; pri_norm = 0.; }
  pri_res /= (pri_norm + OSQP_DIVISION_TOL);
  dua_res /= (dua_norm + OSQP_DIVISION_TOL);
  rho_estimate = work->settings->rho * sqrt(pri_res / dua_res);
  rho_estimate = c_min(c_max(rho_estimate, RHO_MIN), RHO_MAX);
  return rho_estimate;
}

static c_int adapt_rho(OSQPWorkspace *work) {   /* reference src/auxil.c:54-74 */
  c_int exitflag = 0;
  c_float rho_new = compute_rho_estimate(work);
  work->info->rho_estimate = rho_new;
  if (rho_new > work->settings->rho * work->settings->adaptive_rho_tolerance ||
      rho_new < work->settings->rho / work->settings->adaptive_rho_tolerance) {
    exitflag = osqp_update_rho(work, rho_new);
    work->info->rho_updates += 1;
  }
  return exitflag;
}

/* reference src/auxil.c:524-562 */
static void store_solution(OSQPWorkspace *work) {
  c_int i;
  const c_int n = work->data->n, m = work->data->m;
  if (has_solution(work->info)) {
    cuda_api.engine_get_solution(ENGINE(work), work->solution->x, work->solution->y);
  } else {
    const c_int unsc = work->settings->scaling && !work->settings->scaled_termination;
    c_float nrm = 0.;
    for (i = 0; i < n; i++) work->solution->x[i] = OSQP_NAN;
    for (i = 0; i < m; i++) work->solution->y[i] = OSQP_NAN;
    if (work->info->status_val == OSQP_PRIMAL_INFEASIBLE ||
        work->info->status_val == OSQP_PRIMAL_INFEASIBLE_INACCURATE) {
      cuda_api.engine_get_certificates(ENGINE(work), OSQP_NULL, work->delta_y, unsc);
      for (i = 0; i < m; i++) nrm = c_max(nrm, c_absval(work->delta_y[i]));
      for (i = 0; i < m; i++) work->delta_y[i] *= 1. / nrm;
    }
    if (work->info->status_val == OSQP_DUAL_INFEASIBLE ||
        work->info->status_val == OSQP_DUAL_INFEASIBLE_INACCURATE) {
      cuda_api.engine_get_certificates(ENGINE(work), work->delta_x, OSQP_NULL, unsc);
      for (i = 0; i < n; i++) nrm = c_max(nrm, c_absval(work->delta_x[i]));
      for (i = 0; i < n; i++) work->delta_x[i] *= 1. / nrm;
    }
    cold_start(work);
  }
}

/* ------------------------------------------------------------------ polish */
static c_int polish(OSQPWorkspace *work) {   /* reference src/polish.c:212-350 */
  c_int n_low = 0, n_upp = 0, polish_successful;
  osqp_tic(work->timer);
  if (cuda_api.engine_polish(ENGINE(work), work->settings->delta, work->settings->polish_refine_iter,
                             &n_low, &n_upp)) {
    work->info->status_polish = -1;
    return 1;
  }
  work->pol->n_low = n_low;
  work->pol->n_upp = n_upp;
  if (update_info(work, 0, 1, 1)) { work->info->status_polish = -1; return 1; }

  polish_successful = (work->pol->pri_res < work->info->pri_res &&
                       work->pol->dua_res < work->info->dua_res) ||
                      (work->pol->pri_res < work->info->pri_res && work->info->dua_res < 1e-10) ||
                      (work->pol->dua_res < work->info->dua_res && work->info->pri_res < 1e-10);
  if (polish_successful) {
    work->info->obj_val = work->pol->obj_val;
    work->info->pri_res = work->pol->pri_res;
    work->info->dua_res = work->pol->dua_res;
    work->info->status_polish = 1;
    cuda_api.engine_polish_accept(ENGINE(work));
    if (work->settings->verbose) print_polish(work);
  } else {
    work->info->status_polish = -1;
  }
  return 0;
}

/* ------------------------------------------------------------------- solve */
static c_int next_multiple(c_int iter, c_int k) { return (iter / k + 1) * k; }

/* Everything osqp_solve does after its iteration loop (reference src/osqp.c:531-625): last
 * residual check, status, objective, polish, timings, solution.  `iter` is the loop-exit value
 * (iterations run + 1).  Shared with the lockstep batch driver below. */
static c_int solve_finish(OSQPWorkspace *work, c_int iter, c_int can_check_termination, c_int can_print,
                          c_int compute_cost_function, c_int defer_solution) {
  OSQPSettings *s = work->settings;
  OSQPB200Stats st;
  if (!can_check_termination) {
    if (!can_print) {
      if (update_info(work, iter - 1, compute_cost_function, 0)) return 1;
    }
    if (s->verbose && !work->summary_printed) print_summary(work);
    check_termination(work, 0);
  }

  if (!compute_cost_function && has_solution(work->info))
    work->info->obj_val = compute_obj_val(work);

  if (s->verbose && !work->summary_printed) print_summary(work);

  if (work->info->status_val == OSQP_UNSOLVED) {
    if (!check_termination(work, 1)) update_status(work->info, OSQP_MAX_ITER_REACHED);
  }
  if (work->info->status_val == OSQP_TIME_LIMIT_REACHED) {
    if (!check_termination(work, 1)) update_status(work->info, OSQP_TIME_LIMIT_REACHED);
  }

  work->info->rho_estimate = compute_rho_estimate(work);
  work->info->solve_time = osqp_toc(work->timer);

  if (s->polish && work->info->status_val == OSQP_SOLVED) polish(work);

  work->info->run_time = (work->first_run ? work->info->setup_time : work->info->update_time) +
                         work->info->solve_time + work->info->polish_time;
  if (work->first_run) work->first_run = 0;
  work->clear_update_time = 1;
  work->rho_update_from_solve = 0;

  if (s->verbose) print_footer(work->info, s->polish);
  /* lockstep batches read all solutions back together (end of solve_batch_lockstep) */
  if (!(defer_solution && has_solution(work->info))) store_solution(work);
  if (cuda_api.engine_get_stats(ENGINE(work), &st) == 0 && st.pcg_maxed && s->verbose)
    c_print("note: %i linear solves stopped at the CG iteration cap\n", (int)st.pcg_maxed);

  return 0;
}

/* What osqp_solve does before its loop (reference src/osqp.c:288-352) */
static void solve_begin(OSQPWorkspace *work) {
  if (work->clear_update_time == 1) work->info->update_time = 0.0;
  work->rho_update_from_solve = 1;
  osqp_tic(work->timer);
  if (!work->settings->warm_start) cold_start(work);
  cuda_api.engine_begin_solve(ENGINE(work));
}

c_int osqp_solve(OSQPWorkspace *work) {   /* reference src/osqp.c:288-654 */
  c_int exitflag = 0, iter = 0, can_check_termination = 0, can_print, compute_cost_function;
  c_int adapt_now, next, count, poll_clock, chunk;
  c_float temp_run_time;
  OSQPSettings *s;

  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  s = work->settings;
  if (work->clear_update_time == 1) work->info->update_time = 0.0;
  work->rho_update_from_solve = 1;
  can_print = s->verbose;
  compute_cost_function = s->verbose;
  osqp_tic(work->timer);
  if (s->verbose) print_header();
  osqp_start_interrupt_listener();

  if (!s->warm_start) cold_start(work);
  cuda_api.engine_begin_solve(ENGINE(work));

  /* The device runs ahead of the host: iterations are enqueued in runs that end at
   * the next event the host must see (termination check, print line, rho
   * adaptation, clock poll).  check_termination = 0 and no other event => one run. */
  chunk = s->check_termination ? s->check_termination : CHECK_TERMINATION;
  while (iter < s->max_iter) {
    next = s->max_iter;
    if (s->check_termination) next = c_min(next, next_multiple(iter, s->check_termination));
    if (s->verbose) next = c_min(next, iter == 0 ? 1 : next_multiple(iter, PRINT_INTERVAL));
    if (s->adaptive_rho && s->adaptive_rho_interval)
      next = c_min(next, next_multiple(iter, s->adaptive_rho_interval));
    poll_clock = (s->time_limit > 0) || (s->adaptive_rho && !s->adaptive_rho_interval);
    if (poll_clock) next = c_min(next, next_multiple(iter, chunk));
    else next = c_min(next, iter + 1000);          /* SIGINT is polled at least this often */
    count = next - iter;

    can_check_termination = s->check_termination && (next % s->check_termination == 0);
    can_print = s->verbose && ((next % PRINT_INTERVAL == 0) || (next == 1));
    adapt_now = s->adaptive_rho && s->adaptive_rho_interval && (next % s->adaptive_rho_interval == 0);

    if (cuda_api.engine_iterate(ENGINE(work), count, s->alpha,
                                can_check_termination || can_print || adapt_now || next == s->max_iter)) {
      c_eprint("device iteration failed");
      exitflag = 1;
      goto exit;
    }
    iter = next;

    if (osqp_is_interrupted()) {
      update_status(work->info, OSQP_SIGINT);
      c_print("Solver interrupted\n");
      exitflag = 1;
      goto exit;
    }

    if (poll_clock) cuda_api.engine_sync(ENGINE(work));
    if (s->time_limit > 0) {
      temp_run_time = (work->first_run ? work->info->setup_time : work->info->update_time) +
                      osqp_toc(work->timer);
      if (temp_run_time >= s->time_limit) {
        update_status(work->info, OSQP_TIME_LIMIT_REACHED);
        if (s->verbose) c_print("run time limit reached\n");
        can_print = 0;
        can_check_termination = 0;
        iter++;                 /* the loop-exit code below reports iter - 1 */
        goto after_loop;
      }
    }

    if (can_check_termination || can_print) {
      if (update_info(work, iter, compute_cost_function, 0)) { exitflag = 1; goto exit; }
      if (can_print) print_summary(work);
      if (can_check_termination && check_termination(work, 0)) { iter++; goto after_loop; }
    }

    /* automatic adaptive-rho interval: reference src/osqp.c:459-485 */
    if (s->adaptive_rho && !s->adaptive_rho_interval) {
      if (osqp_toc(work->timer) > s->adaptive_rho_fraction * work->info->setup_time) {
        const c_int base = s->check_termination ? s->check_termination : CHECK_TERMINATION;
        s->adaptive_rho_interval = (c_int)((iter + base / 2) / base) * base;   /* round to multiple */
        s->adaptive_rho_interval = c_max(s->adaptive_rho_interval, s->check_termination);
        if (s->adaptive_rho_interval <= 0) s->adaptive_rho_interval = base;
        adapt_now = (iter % s->adaptive_rho_interval == 0);
      }
    }
    if (adapt_now) {
      if (!can_check_termination && !can_print) {
        if (update_info(work, iter, compute_cost_function, 0)) { exitflag = 1; goto exit; }
      }
      if (adapt_rho(work)) {
        c_eprint("Failed rho update");
        exitflag = 1;
        goto exit;
      }
    }
  }
  iter++;   /* mirror the reference's for-loop exit value (max_iter + 1) */

after_loop:
  exitflag = solve_finish(work, iter, can_check_termination, can_print, compute_cost_function, 0);

exit:
  osqp_end_interrupt_listener();
  return exitflag;
}

/* ------------------------------------------------------------ data updates */
static void update_timer_start(OSQPWorkspace *work) {
  if (work->clear_update_time == 1) {
    work->clear_update_time = 0;
    work->info->update_time = 0.0;
  }
  osqp_tic(work->timer);
}

c_int osqp_update_lin_cost(OSQPWorkspace *work, const c_float *q_new) {   /* osqp.c:765-795 */
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  update_timer_start(work);
  if (cuda_api.engine_update_q(ENGINE(work), q_new, work->data->q)) return 1;
  reset_info(work->info);
  work->info->update_time += osqp_toc(work->timer);
  return 0;
}

/* reference src/auxil.c:100-142 through the engine */
static c_int update_rho_vec(OSQPWorkspace *work) {
  c_int changed = 0;
  return cuda_api.engine_retype(ENGINE(work), work->settings->rho, &changed);
}

c_int osqp_update_bounds(OSQPWorkspace *work, const c_float *l_new, const c_float *u_new) {
  c_int i, exitflag;                                                     /* osqp.c:797-848 */
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  update_timer_start(work);
  for (i = 0; i < work->data->m; i++) {
    if (l_new[i] > u_new[i]) {
      c_eprint("lower bound must be lower than or equal to upper bound");
      return 1;
    }
  }
  if (cuda_api.engine_update_bounds(ENGINE(work), l_new, u_new, work->data->l, work->data->u)) return 1;
  reset_info(work->info);
  exitflag = update_rho_vec(work);
  work->info->update_time += osqp_toc(work->timer);
  return exitflag;
}

c_int osqp_update_lower_bound(OSQPWorkspace *work, const c_float *l_new) {   /* osqp.c:850-894 */
  c_int i, exitflag;
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  update_timer_start(work);
  for (i = 0; i < work->data->m; i++) {
    const c_float ls = work->settings->scaling ? work->scaling->E[i] * l_new[i] : l_new[i];
    if (ls > work->data->u[i]) {
      c_eprint("upper bound must be greater than or equal to lower bound");
      return 1;
    }
  }
  if (cuda_api.engine_update_bounds(ENGINE(work), l_new, OSQP_NULL, work->data->l, OSQP_NULL)) return 1;
  reset_info(work->info);
  exitflag = update_rho_vec(work);
  work->info->update_time += osqp_toc(work->timer);
  return exitflag;
}

c_int osqp_update_upper_bound(OSQPWorkspace *work, const c_float *u_new) {   /* osqp.c:896-940 */
  c_int i, exitflag;
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  update_timer_start(work);
  for (i = 0; i < work->data->m; i++) {
    const c_float us = work->settings->scaling ? work->scaling->E[i] * u_new[i] : u_new[i];
    if (us < work->data->l[i]) {
      c_eprint("lower bound must be lower than or equal to upper bound");
      return 1;
    }
  }
  if (cuda_api.engine_update_bounds(ENGINE(work), OSQP_NULL, u_new, OSQP_NULL, work->data->u)) return 1;
  reset_info(work->info);
  exitflag = update_rho_vec(work);
  work->info->update_time += osqp_toc(work->timer);
  return exitflag;
}

c_int osqp_warm_start(OSQPWorkspace *work, const c_float *x, const c_float *y) {   /* osqp.c:942-967 */
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  if (!work->settings->warm_start) work->settings->warm_start = 1;
  return cuda_api.engine_warm_start(ENGINE(work), x, y);
}
c_int osqp_warm_start_x(OSQPWorkspace *work, const c_float *x) {   /* osqp.c:969-988 */
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  if (!work->settings->warm_start) work->settings->warm_start = 1;
  return cuda_api.engine_warm_start(ENGINE(work), x, OSQP_NULL);
}
c_int osqp_warm_start_y(OSQPWorkspace *work, const c_float *y) {   /* osqp.c:990-1007 */
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  if (!work->settings->warm_start) work->settings->warm_start = 1;
  return cuda_api.engine_warm_start(ENGINE(work), OSQP_NULL, y);
}

/* shared tail of osqp_update_P / _A / _P_A: re-equilibrate, refresh the linear system */
static c_int refresh_after_matrix_update(OSQPWorkspace *work) {
  c_int exitflag;
  if (work->settings->scaling) {
    if (cuda_api.engine_scale(ENGINE(work), work->settings->scaling, work->scaling->D,
                              work->scaling->Dinv, work->scaling->E, work->scaling->Einv,
                              &work->scaling->c, work->data->q, work->data->l, work->data->u))
      return 1;
    work->scaling->cinv = 1. / work->scaling->c;
  } else if (cuda_api.engine_scale(ENGINE(work), 0, OSQP_NULL, OSQP_NULL, OSQP_NULL, OSQP_NULL,
                                   OSQP_NULL, OSQP_NULL, OSQP_NULL, OSQP_NULL)) {
    return 1;
  }
  exitflag = cuda_api.engine_refactor(ENGINE(work));
  reset_info(work->info);
  if (exitflag < 0) c_eprint("new KKT matrix is not quasidefinite");
  return exitflag;
}

c_int osqp_update_P(OSQPWorkspace *work, const c_float *Px_new, const c_int *Px_new_idx,
                    c_int P_new_n) {                                       /* osqp.c:1012-1090 */
  c_int exitflag, nnzP;
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  update_timer_start(work);
  nnzP = work->data->P->p[work->data->P->n];
  if (Px_new_idx && P_new_n > nnzP) {
    c_eprint("new number of elements (%i) greater than elements in P (%i)", (int)P_new_n, (int)nnzP);
    return 1;
  }
  if (cuda_api.engine_update_matrices(ENGINE(work), Px_new, Px_new_idx, P_new_n, OSQP_NULL,
                                      OSQP_NULL, 0)) return 1;
  exitflag = refresh_after_matrix_update(work);
  work->info->update_time += osqp_toc(work->timer);
  return exitflag;
}

c_int osqp_update_A(OSQPWorkspace *work, const c_float *Ax_new, const c_int *Ax_new_idx,
                    c_int A_new_n) {                                       /* osqp.c:1093-1170 */
  c_int exitflag, nnzA;
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  update_timer_start(work);
  nnzA = work->data->A->p[work->data->A->n];
  if (Ax_new_idx && A_new_n > nnzA) {
    c_eprint("new number of elements (%i) greater than elements in A (%i)", (int)A_new_n, (int)nnzA);
    return 1;
  }
  if (cuda_api.engine_update_matrices(ENGINE(work), OSQP_NULL, OSQP_NULL, 0, Ax_new, Ax_new_idx,
                                      A_new_n)) return 1;
  exitflag = refresh_after_matrix_update(work);
  work->info->update_time += osqp_toc(work->timer);
  return exitflag;
}

c_int osqp_update_P_A(OSQPWorkspace *work, const c_float *Px_new, const c_int *Px_new_idx,
                      c_int P_new_n, const c_float *Ax_new, const c_int *Ax_new_idx,
                      c_int A_new_n) {                                     /* osqp.c:1173-1279 */
  c_int exitflag, nnzP, nnzA;
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  update_timer_start(work);
  nnzP = work->data->P->p[work->data->P->n];
  nnzA = work->data->A->p[work->data->A->n];
  if (Px_new_idx && P_new_n > nnzP) {
    c_eprint("new number of elements (%i) greater than elements in P (%i)", (int)P_new_n, (int)nnzP);
    return 1;
  }
  if (Ax_new_idx && A_new_n > nnzA) {
    c_eprint("new number of elements (%i) greater than elements in A (%i)", (int)A_new_n, (int)nnzA);
    return 2;
  }
  if (cuda_api.engine_update_matrices(ENGINE(work), Px_new, Px_new_idx, P_new_n, Ax_new,
                                      Ax_new_idx, A_new_n)) return 1;
  exitflag = refresh_after_matrix_update(work);
  work->info->update_time += osqp_toc(work->timer);
  return exitflag;
}

c_int osqp_update_rho(OSQPWorkspace *work, c_float rho_new) {   /* osqp.c:1281-1332 */
  c_int exitflag;
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  if (rho_new <= 0) {
    c_eprint("rho must be positive");
    return 1;
  }
  if (work->rho_update_from_solve == 0) update_timer_start(work);
  work->settings->rho = c_min(c_max(rho_new, RHO_MIN), RHO_MAX);
  exitflag = cuda_api.engine_update_rho(ENGINE(work), work->settings->rho);
  if (work->rho_update_from_solve == 0) work->info->update_time += osqp_toc(work->timer);
  return exitflag;
}

/* -------------------------------------------------------- settings updates
 * reference src/osqp.c:1339-1617: same accepted ranges and messages.          */
#define SETTING_UPDATE(fname, type, field, bad_cond, msg)                \
  c_int fname(OSQPWorkspace *work, type v) {                             \
    if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);         \
    if (bad_cond) { c_eprint(msg); return 1; }                           \
    work->settings->field = v;                                           \
    return 0;                                                            \
  }

SETTING_UPDATE(osqp_update_max_iter, c_int, max_iter, v <= 0, "max_iter must be positive")
SETTING_UPDATE(osqp_update_eps_abs, c_float, eps_abs, v < 0., "eps_abs must be nonnegative")
SETTING_UPDATE(osqp_update_eps_rel, c_float, eps_rel, v < 0., "eps_rel must be nonnegative")
SETTING_UPDATE(osqp_update_eps_prim_inf, c_float, eps_prim_inf, v < 0., "eps_prim_inf must be nonnegative")
SETTING_UPDATE(osqp_update_eps_dual_inf, c_float, eps_dual_inf, v < 0., "eps_dual_inf must be nonnegative")
SETTING_UPDATE(osqp_update_alpha, c_float, alpha, (v <= 0.) || (v >= 2.), "alpha must be between 0 and 2")
SETTING_UPDATE(osqp_update_warm_start, c_int, warm_start, (v != 0) && (v != 1), "warm_start should be either 0 or 1")
SETTING_UPDATE(osqp_update_scaled_termination, c_int, scaled_termination, (v != 0) && (v != 1),
               "scaled_termination should be either 0 or 1")
SETTING_UPDATE(osqp_update_check_termination, c_int, check_termination, v < 0,
               "check_termination should be nonnegative")
SETTING_UPDATE(osqp_update_delta, c_float, delta, v <= 0., "delta must be positive")
SETTING_UPDATE(osqp_update_polish_refine_iter, c_int, polish_refine_iter, v < 0,
               "polish_refine_iter must be nonnegative")
SETTING_UPDATE(osqp_update_verbose, c_int, verbose, (v != 0) && (v != 1), "verbose should be either 0 or 1")
SETTING_UPDATE(osqp_update_time_limit, c_float, time_limit, v < 0., "time_limit must be nonnegative")

c_int osqp_update_polish(OSQPWorkspace *work, c_int polish_new) {   /* osqp.c:1532-1557 */
  if (!work) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  if ((polish_new != 0) && (polish_new != 1)) {
    c_eprint("polish should be either 0 or 1");
    return 1;
  }
  work->settings->polish = polish_new;
  work->info->polish_time = 0.0;
  return 0;
}

/* ---------------------------------------------------------------- additive */
c_int osqp_b200_get_stats(const OSQPWorkspace *work, OSQPB200Stats *out) {
  if (!work || !work->b200 || !ENGINE(work)) return 1;
  return cuda_api.engine_get_stats(ENGINE(work), out);
}
c_int osqp_b200_get_iterates(const OSQPWorkspace *work, c_float *x, c_float *z, c_float *y) {
  if (!work || !work->b200 || !ENGINE(work)) return 1;
  return cuda_api.engine_get_iterates(ENGINE(work), x, z, y);
}
/* handle of the device engine (for bench.py's per-kernel timing through the plugin ABI) */
void *osqp_b200_engine(const OSQPWorkspace *work) {
  if (!work || !work->b200) return OSQP_NULL;
  return (void *)ENGINE(work);
}

/* ------------------------------------------------------------------ batches (additive)
 * BASELINE config 4: many independent small QPs on one GPU.  Every element is an ordinary
 * workspace (own engine, own CUDA stream); `nthreads` host threads pull workspaces from a
 * shared counter and run the ordinary osqp_solve on them, so the device sees up to `nthreads`
 * independent streams of graph launches and the results are exactly those of osqp_solve.
 * Returns the number of workspaces whose osqp_solve returned nonzero. */
#include <pthread.h>
#include <time.h>
typedef struct {
  OSQPWorkspace **works;
  c_int           count;
  volatile c_int  next;
  volatile c_int  failed;
} BatchJob;

static void *batch_worker(void *arg) {
  BatchJob *job = (BatchJob *)arg;
  for (;;) {
    const c_int i = __sync_fetch_and_add(&job->next, 1);
    if (i >= job->count) break;
    if (!job->works[i] || osqp_solve(job->works[i])) __sync_fetch_and_add(&job->failed, 1);
  }
  return OSQP_NULL;
}

/* Lockstep driver for batches of small problems (all engines run the fused single-block
 * kernel, same iteration schedule): ONE host thread walks every problem through the same check
 * intervals; per interval the whole batch costs two kernel launches, one D2H copy and one
 * synchronisation (osqp_cuda_batch_step) instead of ~10 driver calls per problem.  The
 * per-problem logic is osqp_solve's own: update_info_apply / check_termination / adapt_rho /
 * solve_finish, in the same order, so statuses, iteration counts and iterates are those of
 * osqp_solve.  OSQP_B200_LOCKSTEP=0 keeps the threaded path. */
static c_int batch_lockstep_eligible(OSQPWorkspace **works, c_int count) {
  const char *e = getenv("OSQP_B200_LOCKSTEP");
  const OSQPSettings *s0;
  c_int i;
  if ((e && e[0] == '0') || count < 2 || !cuda_api.batch_step || !cuda_api.engine_is_fused) return 0;
  for (i = 0; i < count; i++) {
    const OSQPSettings *s;
    if (!works[i] || !works[i]->b200 || !cuda_api.engine_is_fused(ENGINE(works[i]))) return 0;
    s = works[i]->settings; s0 = works[0]->settings;
    if (s->verbose || s->time_limit > 0 || s->check_termination <= 0) return 0;
    if (s->adaptive_rho && s->adaptive_rho_interval <= 0) return 0;   /* wall-clock rule: per problem */
    if (s->max_iter != s0->max_iter || s->check_termination != s0->check_termination ||
        s->adaptive_rho != s0->adaptive_rho || s->adaptive_rho_interval != s0->adaptive_rho_interval ||
        s->alpha != s0->alpha) return 0;
  }
  return 1;
}

static c_int solve_batch_lockstep(OSQPWorkspace **works, c_int count) {
  const OSQPSettings *s = works[0]->settings;
  OSQPWorkspace **act = (OSQPWorkspace **)malloc(sizeof(*act) * (size_t)count);
  osqp_cuda_engine **eng = (osqp_cuda_engine **)malloc(sizeof(*eng) * (size_t)count);
  osqp_cuda_resid **res = (osqp_cuda_resid **)malloc(sizeof(*res) * (size_t)count);
  c_int nact = count, failed = 0, iter = 0, i, k, can_check = 0;
  if (!act || !eng || !res) { free(act); free(eng); free(res); return count; }
  /* OSQP_B200_TIMING: where a batch spends its time (host clock; the device work is inside batch_step) */
  const int timing = getenv("OSQP_B200_TIMING") != OSQP_NULL;
  struct timespec ts0, ts1;
  double t_begin = 0.0, t_step = 0.0, t_host = 0.0, t_read = 0.0;
#define LS_TIC() do { if (timing) clock_gettime(CLOCK_MONOTONIC, &ts0); } while (0)
#define LS_TOC(acc) do { if (timing) { clock_gettime(CLOCK_MONOTONIC, &ts1); \
    acc += (double)(ts1.tv_sec - ts0.tv_sec) + 1e-9 * (double)(ts1.tv_nsec - ts0.tv_nsec); } } while (0)
  osqp_start_interrupt_listener();
  LS_TIC();
  if (cuda_api.batch_begin_solve) {
    /* solve_begin for the whole batch: the host part per problem, the device part (cold start, CG
     * statistics reset) in one launch for the cold problems and one for the warm-started ones */
    c_int pass;
    for (i = 0; i < count; i++) {
      OSQPWorkspace *w = works[i];
      act[i] = w;
      if (w->clear_update_time == 1) w->info->update_time = 0.0;
      w->rho_update_from_solve = 1;
      osqp_tic(w->timer);
    }
    for (pass = 0; pass < 2; pass++) {          /* pass 0: cold, pass 1: warm-started */
      for (i = 0, k = 0; i < count; i++)
        if ((works[i]->settings->warm_start ? 1 : 0) == pass) eng[k++] = ENGINE(works[i]);
      if (k && cuda_api.batch_begin_solve(eng, k, pass == 0)) {
        c_eprint("device batch start failed");
        osqp_end_interrupt_listener();
        free(act); free(eng); free(res);
        return count;
      }
    }
  } else {
    for (i = 0; i < count; i++) { act[i] = works[i]; solve_begin(works[i]); }
  }
  LS_TOC(t_begin);

  while (nact > 0 && iter < s->max_iter) {
    c_int next = s->max_iter, adapt_now, store, do_resid;
    next = c_min(next, next_multiple(iter, s->check_termination));
    if (s->adaptive_rho) next = c_min(next, next_multiple(iter, s->adaptive_rho_interval));
    next = c_min(next, iter + 1000);
    can_check = (next % s->check_termination == 0);
    adapt_now = s->adaptive_rho && (next % s->adaptive_rho_interval == 0);
    do_resid = can_check || adapt_now;
    store = do_resid || next == s->max_iter;
    for (i = 0; i < nact; i++) { eng[i] = ENGINE(act[i]); res[i] = &PRIV(act[i])->res; }
    LS_TIC();
    k = cuda_api.batch_step(eng, nact, next - iter, s->alpha, store, do_resid, res);
    LS_TOC(t_step);
    if (k || osqp_is_interrupted()) {
      c_eprint("device batch step failed or interrupted");
      failed += nact;
      nact = 0;
      break;
    }
    iter = next;
    LS_TIC();
    for (i = 0, k = 0; i < nact; i++) {
      OSQPWorkspace *w = act[i];
      c_int done = 0;
      if (do_resid) update_info_apply(w, iter, 0, 0);
      if (can_check && check_termination(w, 0)) done = 1;
      if (!done && adapt_now && adapt_rho(w)) { c_eprint("Failed rho update"); failed++; continue; }
      if (done) { if (solve_finish(w, iter + 1, 1, 0, 0, 1)) failed++; }
      else act[k++] = w;
    }
    LS_TOC(t_host);
    nact = k;
  }
  for (i = 0; i < nact; i++)          /* iteration limit reached */
    if (solve_finish(act[i], iter + 1, can_check, 0, 0, 1)) failed++;
  /* solutions of every problem that has one, read back together */
  {
    c_float **xs = (c_float **)malloc(sizeof(*xs) * (size_t)count), **ys = (c_float **)malloc(sizeof(*ys) * (size_t)count);
    k = 0;
    if (xs && ys) {
      for (i = 0; i < count; i++)
        if (has_solution(works[i]->info)) {
          eng[k] = ENGINE(works[i]); xs[k] = works[i]->solution->x; ys[k] = works[i]->solution->y; k++;
        }
      LS_TIC();
      if (k && cuda_api.batch_get_solutions(eng, k, xs, ys)) failed += k;
      LS_TOC(t_read);
    } else failed = count;
    free(xs); free(ys);
  }
  if (timing)
    c_print("[timing] lockstep batch of %d: begin %.1f ms, device steps %.1f ms, per-problem host work %.1f ms, read-back %.1f ms\n",
            (int)count, 1e3 * t_begin, 1e3 * t_step, 1e3 * t_host, 1e3 * t_read);
#undef LS_TIC
#undef LS_TOC
  osqp_end_interrupt_listener();
  free(act); free(eng); free(res);
  return failed;
}

c_int osqp_solve_batch(OSQPWorkspace **works, c_int count, c_int nthreads) {
  BatchJob  job;
  pthread_t tid[256];
  c_int     t, started = 0;
  if (!works || count < 0) return osqp_error(OSQP_WORKSPACE_NOT_INIT_ERROR);
  if (nthreads < 1) nthreads = 1;
  if (nthreads > 256) nthreads = 256;
  if (nthreads > count) nthreads = count;
  if (batch_lockstep_eligible(works, count)) return solve_batch_lockstep(works, count);
  job.works = works; job.count = count; job.next = 0; job.failed = 0;
  for (t = 1; t < nthreads; t++) {
    if (pthread_create(&tid[started], OSQP_NULL, batch_worker, &job)) continue;
    started++;
  }
  batch_worker(&job);                       /* the calling thread works too */
  for (t = 0; t < started; t++) pthread_join(tid[t], OSQP_NULL);
  return job.failed;
}
