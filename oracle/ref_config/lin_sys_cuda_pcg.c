/* Test infrastructure (oracle/): the reference's linear-system dispatcher with the cuda_pcg backend
 * wired in, exactly the patch INTEGRATION.md section A describes.  `make -C oracle ref_cuda` compiles
 * the UNMODIFIED reference sources with this file in place of src/lin_sys.c (plus the enum value of
 * include/constants.h:35 as -DCUDA_PCG_SOLVER=2 and the check of src/auxil.c:881-891 edited on a
 * temporary copy) into oracle/_ref/libosqp_ref_cuda.so.  tests/test_gpu_dropin.py then runs the
 * reference's OWN osqp_solve loop (src/osqp.c:247-625) with libosqp_cuda_pcg.so as its KKT solver.
 *
 * Follows the structure of the reference's src/lin_sys.c:15-75 (load / unload / init switch) and
 * lin_sys/direct/pardiso/pardiso_loader.c:63-102 (dlopen through lib_handler.c), written from scratch. */
#include "lin_sys.h"
#include "lib_handler.h"
#include "qdldl_interface.h"
#include <stdlib.h>

/* defined in lin_sys/lib_handler.c:66 but not declared in lib_handler.h (pardiso_loader.c declares it too) */
void *lh_load_sym(soHandle_t h, const char *symName);

const char *LINSYS_SOLVER_NAME[] = { "qdldl", "mkl pardiso", "cuda pcg" };

typedef c_int (*cuda_pcg_init_fn)(LinSysSolver **s, const csc *P, const csc *A, c_float sigma,
                                  const c_float *rho_vec, c_int polish);
typedef c_int (*cuda_pcg_probe_fn)(void);

static soHandle_t       cuda_handle = OSQP_NULL;
static cuda_pcg_init_fn cuda_init   = OSQP_NULL;
static int              cuda_users  = 0;      /* workspaces (and their polish instances) share it */

static c_int cuda_pcg_load(void) {
  const char *path = getenv("OSQP_CUDA_PCG_LIB");
  cuda_pcg_probe_fn probe;
  if (cuda_users > 0) { cuda_users++; return 0; }
  cuda_handle = lh_load_lib(path ? path : "libosqp_cuda_pcg.so");
  if (!cuda_handle) return 1;
  cuda_init = (cuda_pcg_init_fn)lh_load_sym(cuda_handle, "init_linsys_solver_cuda_pcg");
  probe     = (cuda_pcg_probe_fn)lh_load_sym(cuda_handle, "osqp_cuda_probe");
  if (!cuda_init || !probe || probe()) {     /* no usable sm_100 device: fail loudly, no fallback */
    lh_unload_lib(cuda_handle);
    cuda_handle = OSQP_NULL; cuda_init = OSQP_NULL;
    return 1;
  }
  cuda_users = 1;
  return 0;
}

/* Only the user count drops: the library stays mapped.  Its statically linked CUDA runtime does not
 * survive a dlclose + dlopen cycle (a second osqp_setup after osqp_cleanup would crash or leak the
 * context); the product's own loader (osqp_b200/csrc/lin_sys_host.c) never unloads either.        */
static c_int cuda_pcg_unload(void) {
  if (cuda_users > 0) --cuda_users;
  return 0;
}

c_int load_linsys_solver(enum linsys_solver_type linsys_solver) {
  if (linsys_solver == CUDA_PCG_SOLVER) return cuda_pcg_load();
  return 0;                                   /* QDLDL is compiled in */
}

c_int unload_linsys_solver(enum linsys_solver_type linsys_solver) {
  if (linsys_solver == CUDA_PCG_SOLVER) return cuda_pcg_unload();
  return 0;
}

c_int init_linsys_solver(LinSysSolver **s, const csc *P, const csc *A, c_float sigma,
                         const c_float *rho_vec, enum linsys_solver_type linsys_solver, c_int polish) {
  if (linsys_solver == CUDA_PCG_SOLVER) {
    if (!cuda_init) return OSQP_LINSYS_SOLVER_LOAD_ERROR;
    return cuda_init(s, P, A, sigma, rho_vec, polish);
  }
  return init_linsys_solver_qdldl((qdldl_solver **)s, P, A, sigma, rho_vec, polish);
}
