/* lin_sys_host.c -- the linear-system plugin boundary on the host.
 *
 * Mirrors reference src/lin_sys.c:15-75 (load / unload / init dispatch) and the
 * shared-library handler of reference lin_sys/lib_handler.c:7-155 +
 * lin_sys/direct/pardiso/pardiso_loader.c:63-102: the CUDA backend lives in its
 * own shared object, opened with dlopen() and resolved symbol by symbol into a
 * static function table.  Unlike the Pardiso loader this one is ref-counted, so
 * several live workspaces (and the polish instance) can share it.             */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <ctype.h>
#include "osqp_host.h"

const char *LINSYS_SOLVER_NAME[] = {"qdldl", "mkl pardiso", "cuda pcg"};

cuda_pcg_api cuda_api;
static void *cuda_pcg_handle = OSQP_NULL;
static int   cuda_pcg_refs = 0;
static char  plugin_path_override[4096] = "";

void osqp_b200_set_plugin_path(const char *path) {
  if (!path) { plugin_path_override[0] = '\0'; return; }
  strncpy(plugin_path_override, path, sizeof(plugin_path_override) - 1);
  plugin_path_override[sizeof(plugin_path_override) - 1] = '\0';
}

/* ---- generic handler: open / close / symbol lookup with the name variants the
 *      reference tries (as given, lower_, UPPER_, given_, lower, UPPER)          */
static void *lh_load_lib(const char *name) {
  void *h;
  if (!name) { c_eprint("no library name given"); return OSQP_NULL; }
  h = dlopen(name, RTLD_LAZY);
  if (!h) c_eprint("Error while loading dynamic library %s: %s", name, dlerror());
  return h;
}
static c_int lh_unload_lib(void *h) { return (c_int)dlclose(h); }

static void *lh_load_sym(void *h, const char *sym) {
  char buf[260];
  size_t len = strlen(sym), k;
  int variant;
  if (len > 256) return OSQP_NULL;
  for (variant = 0; variant < 6; variant++) {
    void *s;
    for (k = 0; k < len; k++) {
      int ch = (unsigned char)sym[k];
      if (variant == 1 || variant == 4) ch = tolower(ch);
      if (variant == 2 || variant == 5) ch = toupper(ch);
      buf[k] = (char)ch;
    }
    if (variant >= 1 && variant <= 3) buf[k++] = '_';
    buf[k] = '\0';
    dlerror();
    s = dlsym(h, buf);
    if (!dlerror() && s) return s;
  }
  c_eprint("Cannot find symbol %s in dynamic library", sym);
  return OSQP_NULL;
}

/* default location: the directory this library was loaded from */
static void default_plugin_path(char *out, size_t cap) {
  Dl_info info;
  const char *env = getenv("OSQP_CUDA_PCG_LIB");
  if (plugin_path_override[0]) { snprintf(out, cap, "%s", plugin_path_override); return; }
  if (env && env[0]) { snprintf(out, cap, "%s", env); return; }
  if (dladdr((void *)&default_plugin_path, &info) && info.dli_fname) {
    const char *slash = strrchr(info.dli_fname, '/');
    if (slash) {
      snprintf(out, cap, "%.*s/libosqp_cuda_pcg.so", (int)(slash - info.dli_fname), info.dli_fname);
      return;
    }
  }
  snprintf(out, cap, "libosqp_cuda_pcg.so");
}

#define RESOLVE(field, name)                                            \
  do {                                                                  \
    *(void **)(&cuda_api.field) = lh_load_sym(cuda_pcg_handle, name);   \
    if (!cuda_api.field) goto fail;                                     \
  } while (0)

c_int lh_load_cuda_pcg(const char *libname) {
  char path[4096];
  if (cuda_pcg_handle) { cuda_pcg_refs++; return 0; }
  if (libname) snprintf(path, sizeof(path), "%s", libname);
  else default_plugin_path(path, sizeof(path));
  cuda_pcg_handle = lh_load_lib(path);
  if (!cuda_pcg_handle) return 1;
  RESOLVE(init_linsys, "init_linsys_solver_cuda_pcg");
  RESOLVE(probe, "osqp_cuda_probe");
  RESOLVE(engine_create, "osqp_cuda_engine_create");
  RESOLVE(engine_free, "osqp_cuda_engine_free");
  RESOLVE(engine_scale, "osqp_cuda_engine_scale");
  RESOLVE(engine_init_linsys, "osqp_cuda_engine_init_linsys");
  RESOLVE(engine_update_rho, "osqp_cuda_engine_update_rho");
  RESOLVE(engine_retype, "osqp_cuda_engine_retype_constraints");
  RESOLVE(engine_iterate, "osqp_cuda_engine_iterate");
  RESOLVE(engine_residuals, "osqp_cuda_engine_residuals");
  RESOLVE(engine_cold_start, "osqp_cuda_engine_cold_start");
  RESOLVE(engine_begin_solve, "osqp_cuda_engine_begin_solve");
  RESOLVE(engine_warm_start, "osqp_cuda_engine_warm_start");
  RESOLVE(engine_update_q, "osqp_cuda_engine_update_q");
  RESOLVE(engine_update_bounds, "osqp_cuda_engine_update_bounds");
  RESOLVE(engine_update_matrices, "osqp_cuda_engine_update_matrices");
  RESOLVE(engine_refactor, "osqp_cuda_engine_refactor");
  RESOLVE(engine_get_solution, "osqp_cuda_engine_get_solution");
  RESOLVE(engine_get_iterates, "osqp_cuda_engine_get_iterates");
  RESOLVE(engine_get_certificates, "osqp_cuda_engine_get_certificates");
  RESOLVE(engine_polish, "osqp_cuda_engine_polish");
  RESOLVE(engine_polish_accept, "osqp_cuda_engine_polish_accept");
  RESOLVE(engine_get_stats, "osqp_cuda_engine_get_stats");
  RESOLVE(engine_sync, "osqp_cuda_engine_sync");
  RESOLVE(engine_is_fused, "osqp_cuda_engine_is_fused");
  RESOLVE(batch_step, "osqp_cuda_batch_step");
  RESOLVE(batch_get_solutions, "osqp_cuda_batch_get_solutions");
  RESOLVE(batch_begin_solve, "osqp_cuda_batch_begin_solve");
  /* no CPU fallback: a box without a usable B200 fails here, loudly */
  if (cuda_api.probe()) {
    c_eprint("libosqp_cuda_pcg.so loaded but no usable sm_100 CUDA device was found");
    goto fail;
  }
  cuda_pcg_refs = 1;
  return 0;
fail:
  lh_unload_lib(cuda_pcg_handle);
  cuda_pcg_handle = OSQP_NULL;
  memset(&cuda_api, 0, sizeof(cuda_api));
  return 1;
}

c_int lh_unload_cuda_pcg(void) {
  /* Only the count drops.  The object stays mapped: the CUDA runtime inside it
   * owns process-wide state (context, module images) that does not survive
   * dlclose + dlopen cycles, and the function table must stay valid for the
   * other live workspaces (cf. the note in pardiso_loader.c:96-100).          */
  if (cuda_pcg_refs > 0) cuda_pcg_refs--;
  return 0;
}
c_int lh_cuda_pcg_loaded(void) { return cuda_pcg_handle != OSQP_NULL; }

/* ---- dispatcher: reference src/lin_sys.c:15-75 ---- */
c_int load_linsys_solver(enum linsys_solver_type linsys_solver) {
  switch (linsys_solver) {
  case CUDA_PCG_SOLVER:
    return lh_load_cuda_pcg(OSQP_NULL);
  default:
    c_eprint("linear system solver %d is not part of this library (only cuda pcg)", (int)linsys_solver);
    return 1;
  }
}

c_int unload_linsys_solver(enum linsys_solver_type linsys_solver) {
  switch (linsys_solver) {
  case CUDA_PCG_SOLVER:
    return lh_unload_cuda_pcg();
  default:
    return 0;
  }
}

c_int init_linsys_solver(LinSysSolver **s, const csc *P, const csc *A, c_float sigma,
                         const c_float *rho_vec, enum linsys_solver_type linsys_solver,
                         c_int polish) {
  switch (linsys_solver) {
  case CUDA_PCG_SOLVER:
    if (!cuda_pcg_handle) { *s = OSQP_NULL; return OSQP_LINSYS_SOLVER_INIT_ERROR; }
    return cuda_api.init_linsys((cuda_pcg_solver **)s, P, A, sigma, rho_vec, polish);
  default:
    *s = OSQP_NULL;
    return OSQP_LINSYS_SOLVER_INIT_ERROR;
  }
}
