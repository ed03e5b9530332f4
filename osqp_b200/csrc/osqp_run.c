/* osqp_run -- solve a QP stored on disk through the OSQP C API, without Python.
 *
 * the reference only has the DDEBUG dump helpers (src/util.c:393-472,
 * dump_csc_matrix / dump_vec, text, one file per object); this is one binary file per problem
 * and a runner, so that a benchmark instance can be reproduced from C alone.
 *
 * The library is dlopen()ed and only the reference's own entry points are used
 * (osqp_set_default_settings / osqp_setup / osqp_solve / osqp_cleanup, include/osqp.h), so the
 * same binary drives libosqp_b200.so (linsys_solver 2 = cuda pcg) and a build of the reference
 * (linsys_solver 0 = qdldl): the ABI is the drop-in boundary.
 *
 *   osqp_run problem.bin [--lib path/libosqp_b200.so] [--solver 2] [--eps 1e-3] [--max-iter N]
 *            [--rho-interval 50] [--polish 0|1] [--scaling 10] [--verbose 0|1] [--out solution.bin]
 *            [--info]      (print the header of the file and exit: no library, no GPU needed)
 *
 * File format (little endian): 8 bytes "OSQPB200", int64 n, m, nnzP, nnzA, then
 *   P.p[n+1] P.i[nnzP] (int64) P.x[nnzP] (f64)   upper-triangular CSC
 *   A.p[n+1] A.i[nnzA] (int64) A.x[nnzA] (f64)   CSC
 *   q[n] l[m] u[m] (f64)
 * Solution file: 8 bytes "OSQPSOL0", int64 n, m, status_val, iter, f64 obj_val, pri_res, dua_res,
 *   x[n], y[m].
 */
#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#include "osqp.h"

static int rd(FILE *f, void *p, size_t bytes) { return bytes && fread(p, 1, bytes, f) != bytes; }

int main(int argc, char **argv) {
  const char *path = NULL, *lib = NULL, *out = NULL;
  long long hdr[4];
  char magic[8];
  int i, info_only = 0, solver = 2, polish = 0, scaling = -1, verbose = 0;
  long long max_iter = -1, rho_interval = 50;
  double eps = 1e-3;
  FILE *f;

  for (i = 1; i < argc; i++) {
    if (!strcmp(argv[i], "--info")) info_only = 1;
    else if (i + 1 < argc && !strcmp(argv[i], "--lib")) lib = argv[++i];
    else if (i + 1 < argc && !strcmp(argv[i], "--out")) out = argv[++i];
    else if (i + 1 < argc && !strcmp(argv[i], "--solver")) solver = atoi(argv[++i]);
    else if (i + 1 < argc && !strcmp(argv[i], "--eps")) eps = atof(argv[++i]);
    else if (i + 1 < argc && !strcmp(argv[i], "--max-iter")) max_iter = atoll(argv[++i]);
    else if (i + 1 < argc && !strcmp(argv[i], "--rho-interval")) rho_interval = atoll(argv[++i]);
    else if (i + 1 < argc && !strcmp(argv[i], "--polish")) polish = atoi(argv[++i]);
    else if (i + 1 < argc && !strcmp(argv[i], "--scaling")) scaling = atoi(argv[++i]);
    else if (i + 1 < argc && !strcmp(argv[i], "--verbose")) verbose = atoi(argv[++i]);
    else if (argv[i][0] != '-' && !path) path = argv[i];
    else { fprintf(stderr, "osqp_run: unknown argument %s\n", argv[i]); return 2; }
  }
  if (!path) { fprintf(stderr, "usage: osqp_run problem.bin [--lib lib.so] [--solver N] [--eps e] ...\n"); return 2; }
  f = fopen(path, "rb");
  if (!f || rd(f, magic, 8) || memcmp(magic, "OSQPB200", 8) || rd(f, hdr, sizeof(hdr))) {
    fprintf(stderr, "osqp_run: %s is not an OSQPB200 problem file\n", path);
    return 2;
  }
  {
    const long long n = hdr[0], m = hdr[1], nnzP = hdr[2], nnzA = hdr[3];
    c_int *Pp, *Pi, *Ap, *Ai;
    c_float *Px, *Ax, *q, *l, *u;
    if (n <= 0 || m < 0 || nnzP < 0 || nnzA < 0) { fprintf(stderr, "osqp_run: bad header\n"); return 2; }
    printf("{\"file\": \"%s\", \"n\": %lld, \"m\": %lld, \"nnzP\": %lld, \"nnzA\": %lld}\n", path, n, m, nnzP, nnzA);
    if (info_only) { fclose(f); return 0; }
    Pp = (c_int *)malloc(sizeof(c_int) * (size_t)(n + 1)); Pi = (c_int *)malloc(sizeof(c_int) * (size_t)(nnzP + 1));
    Px = (c_float *)malloc(sizeof(c_float) * (size_t)(nnzP + 1));
    Ap = (c_int *)malloc(sizeof(c_int) * (size_t)(n + 1)); Ai = (c_int *)malloc(sizeof(c_int) * (size_t)(nnzA + 1));
    Ax = (c_float *)malloc(sizeof(c_float) * (size_t)(nnzA + 1));
    q = (c_float *)malloc(sizeof(c_float) * (size_t)n);
    l = (c_float *)malloc(sizeof(c_float) * (size_t)(m + 1)); u = (c_float *)malloc(sizeof(c_float) * (size_t)(m + 1));
    if (!Pp || !Pi || !Px || !Ap || !Ai || !Ax || !q || !l || !u) { fprintf(stderr, "osqp_run: out of memory\n"); return 2; }
    if (rd(f, Pp, 8 * (size_t)(n + 1)) || rd(f, Pi, 8 * (size_t)nnzP) || rd(f, Px, 8 * (size_t)nnzP) ||
        rd(f, Ap, 8 * (size_t)(n + 1)) || rd(f, Ai, 8 * (size_t)nnzA) || rd(f, Ax, 8 * (size_t)nnzA) ||
        rd(f, q, 8 * (size_t)n) || rd(f, l, 8 * (size_t)m) || rd(f, u, 8 * (size_t)m)) {
      fprintf(stderr, "osqp_run: %s is truncated\n", path);
      return 2;
    }
    fclose(f);
    if (Pp[n] != nnzP || Ap[n] != nnzA) { fprintf(stderr, "osqp_run: column pointers do not match the header\n"); return 2; }
    {
      /* the file is untrusted: monotone column pointers, row indices in range (these arrays reach
       * device kernels as they are) */
      long long j, e;
      int bad = Pp[0] != 0 || Ap[0] != 0;
      for (j = 0; j < n && !bad; ++j) bad = Pp[j + 1] < Pp[j] || Ap[j + 1] < Ap[j];
      for (e = 0; e < nnzP && !bad; ++e) bad = Pi[e] < 0 || Pi[e] >= n;
      for (e = 0; e < nnzA && !bad; ++e) bad = Ai[e] < 0 || Ai[e] >= m;
      if (bad) { fprintf(stderr, "osqp_run: %s holds column pointers or row indices out of range\n", path); return 2; }
    }
    {
      /* the reference's public entry points, resolved from whichever library was asked for */
      char defpath[4096];
      void *h;
      void (*set_defaults)(OSQPSettings *);
      c_int (*setup)(OSQPWorkspace **, const OSQPData *, const OSQPSettings *);
      c_int (*solve)(OSQPWorkspace *);
      c_int (*cleanup)(OSQPWorkspace *);
      OSQPSettings settings;
      OSQPData data;
      OSQPWorkspace *work = NULL;
      csc P, A;
      c_int flag;
      if (!lib) {      /* default: libosqp_b200.so next to this executable */
        ssize_t k = readlink("/proc/self/exe", defpath, sizeof(defpath) - 32);
        if (k <= 0) { fprintf(stderr, "osqp_run: pass --lib\n"); return 2; }
        defpath[k] = 0;
        { char *s = strrchr(defpath, '/'); if (s) s[1] = 0; }
        strcat(defpath, "libosqp_b200.so");
        lib = defpath;
      }
      h = dlopen(lib, RTLD_NOW | RTLD_LOCAL);
      if (!h) { fprintf(stderr, "osqp_run: cannot load %s: %s\n", lib, dlerror()); return 3; }
      *(void **)&set_defaults = dlsym(h, "osqp_set_default_settings");
      *(void **)&setup = dlsym(h, "osqp_setup");
      *(void **)&solve = dlsym(h, "osqp_solve");
      *(void **)&cleanup = dlsym(h, "osqp_cleanup");
      if (!set_defaults || !setup || !solve || !cleanup) { fprintf(stderr, "osqp_run: %s is not an OSQP library\n", lib); return 3; }
      P.nzmax = nnzP; P.m = n; P.n = n; P.p = Pp; P.i = Pi; P.x = Px; P.nz = -1;
      A.nzmax = nnzA; A.m = m; A.n = n; A.p = Ap; A.i = Ai; A.x = Ax; A.nz = -1;
      data.n = n; data.m = m; data.P = &P; data.A = &A; data.q = q; data.l = l; data.u = u;
      set_defaults(&settings);
      settings.linsys_solver = (enum linsys_solver_type)solver;
      settings.eps_abs = settings.eps_rel = eps;
      settings.adaptive_rho_interval = rho_interval;
      settings.polish = polish;
      settings.verbose = verbose;
      if (max_iter > 0) settings.max_iter = max_iter;
      if (scaling >= 0) settings.scaling = scaling;
      flag = setup(&work, &data, &settings);
      if (flag || !work) { fprintf(stderr, "osqp_run: osqp_setup failed (%lld)\n", (long long)flag); return 4; }
      flag = solve(work);
      if (flag) { fprintf(stderr, "osqp_run: osqp_solve failed (%lld)\n", (long long)flag); cleanup(work); return 5; }
      printf("{\"status\": \"%s\", \"status_val\": %lld, \"iter\": %lld, \"obj_val\": %.15g, \"pri_res\": %.6g, "
             "\"dua_res\": %.6g, \"setup_time\": %.6f, \"solve_time\": %.6f, \"rho_updates\": %lld}\n",
             work->info->status, (long long)work->info->status_val, (long long)work->info->iter,
             work->info->obj_val, work->info->pri_res, work->info->dua_res, work->info->setup_time,
             work->info->solve_time, (long long)work->info->rho_updates);
      if (out) {
        FILE *g = fopen(out, "wb");
        long long sh[4] = {n, m, (long long)work->info->status_val, (long long)work->info->iter};
        double sd[3] = {work->info->obj_val, work->info->pri_res, work->info->dua_res};
        if (!g || fwrite("OSQPSOL0", 1, 8, g) != 8 || fwrite(sh, 8, 4, g) != 4 || fwrite(sd, 8, 3, g) != 3 ||
            fwrite(work->solution->x, 8, (size_t)n, g) != (size_t)n ||
            (m && fwrite(work->solution->y, 8, (size_t)m, g) != (size_t)m)) {
          fprintf(stderr, "osqp_run: cannot write %s\n", out);
          return 6;
        }
        fclose(g);
      }
      cleanup(work);
    }
    free(Pp); free(Pi); free(Px); free(Ap); free(Ai); free(Ax); free(q); free(l); free(u);
  }
  return 0;
}
