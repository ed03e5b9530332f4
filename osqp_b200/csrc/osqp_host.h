/* osqp_host.h -- internals of libosqp_b200.so (plain C host side).
 * Nothing here is CUDA: the device work is reached through the function table
 * that lh_load_cuda_pcg() fills from libosqp_cuda_pcg.so with dlopen/dlsym.   */
#ifndef OSQP_HOST_H
#define OSQP_HOST_H

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <time.h>
#include "osqp_cuda_pcg.h"

#define c_print printf
#define c_eprint(...) do { printf("ERROR in %s: ", __FUNCTION__); printf(__VA_ARGS__); printf("\n"); } while (0)
#define c_max(a, b) (((a) > (b)) ? (a) : (b))
#define c_min(a, b) (((a) < (b)) ? (a) : (b))
#define c_absval(x) (((x) < 0) ? -(x) : (x))

/* reference include/util.h:129-132 (Linux branch) */
struct OSQP_TIMER {
  struct timespec tic;
  struct timespec toc;
};
void    osqp_tic(OSQPTimer *t);
c_float osqp_toc(OSQPTimer *t);

/* ---- function table of the CUDA plugin (filled by lh_load_cuda_pcg) ---- */
typedef struct {
  c_int (*init_linsys)(cuda_pcg_solver **, const csc *, const csc *, c_float, const c_float *, c_int);
  c_int (*probe)(void);
  c_int (*engine_create)(osqp_cuda_engine **, c_int, c_int, const csc *, const csc *,
                         const c_float *, const c_float *, const c_float *);
  void  (*engine_free)(osqp_cuda_engine *);
  c_int (*engine_scale)(osqp_cuda_engine *, c_int, c_float *, c_float *, c_float *, c_float *,
                        c_float *, c_float *, c_float *, c_float *);
  c_int (*engine_init_linsys)(osqp_cuda_engine *, c_float, c_float, cuda_pcg_solver **);
  c_int (*engine_update_rho)(osqp_cuda_engine *, c_float);
  c_int (*engine_retype)(osqp_cuda_engine *, c_float, c_int *);
  c_int (*engine_iterate)(osqp_cuda_engine *, c_int, c_float, c_int);
  c_int (*engine_residuals)(osqp_cuda_engine *, osqp_cuda_resid *, c_int);
  c_int (*engine_cold_start)(osqp_cuda_engine *);
  c_int (*engine_begin_solve)(osqp_cuda_engine *);
  c_int (*engine_warm_start)(osqp_cuda_engine *, const c_float *, const c_float *);
  c_int (*engine_update_q)(osqp_cuda_engine *, const c_float *, c_float *);
  c_int (*engine_update_bounds)(osqp_cuda_engine *, const c_float *, const c_float *, c_float *, c_float *);
  c_int (*engine_update_matrices)(osqp_cuda_engine *, const c_float *, const c_int *, c_int,
                                  const c_float *, const c_int *, c_int);
  c_int (*engine_refactor)(osqp_cuda_engine *);
  c_int (*engine_get_solution)(osqp_cuda_engine *, c_float *, c_float *);
  c_int (*engine_get_iterates)(osqp_cuda_engine *, c_float *, c_float *, c_float *);
  c_int (*engine_get_certificates)(osqp_cuda_engine *, c_float *, c_float *, c_int);
  c_int (*engine_polish)(osqp_cuda_engine *, c_float, c_int, c_int *, c_int *);
  c_int (*engine_polish_accept)(osqp_cuda_engine *);
  c_int (*engine_get_stats)(const osqp_cuda_engine *, OSQPB200Stats *);
  c_int (*engine_sync)(osqp_cuda_engine *);
  c_int (*engine_is_fused)(const osqp_cuda_engine *);
  c_int (*batch_step)(osqp_cuda_engine **, c_int, c_int, c_float, c_int, c_int, osqp_cuda_resid **);
  c_int (*batch_get_solutions)(osqp_cuda_engine **, c_int, c_float **, c_float **);
  c_int (*batch_begin_solve)(osqp_cuda_engine **, c_int, c_int);
} cuda_pcg_api;

extern cuda_pcg_api cuda_api;       /* valid while the plugin is loaded            */
c_int lh_load_cuda_pcg(const char *libname);   /* 0 ok, 1 failure (ref-counted)   */
c_int lh_unload_cuda_pcg(void);
c_int lh_cuda_pcg_loaded(void);

/* ---- private part of the workspace (OSQPWorkspace::b200) ---- */
typedef struct {
  osqp_cuda_engine *engine;
  osqp_cuda_resid   res;         /* scalars of the last residual pass             */
  c_int             res_valid;
  c_int             plugin_loaded;
  c_int             pcg_noncvx;  /* negative curvature reported by the device CG  */
} OSQPB200Priv;

/* ---- helpers shared between the host files ---- */
c_int  osqp_error_(enum osqp_error_type code, const char *function_name);
#define osqp_error(code) osqp_error_(code, __FUNCTION__)
c_int  validate_data(const OSQPData *data);
c_int  validate_settings(const OSQPSettings *settings);
void   update_status(OSQPInfo *info, c_int status_val);
void   reset_info(OSQPInfo *info);
c_int  has_solution(const OSQPInfo *info);
void   print_setup_header(const OSQPWorkspace *work);
void   print_header(void);
void   print_summary(OSQPWorkspace *work);
void   print_polish(OSQPWorkspace *work);
void   print_footer(const OSQPInfo *info, c_int polish);
void   osqp_start_interrupt_listener(void);
void   osqp_end_interrupt_listener(void);
int    osqp_is_interrupted(void);

#endif /* OSQP_HOST_H */
