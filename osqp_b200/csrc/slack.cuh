// slack.cuh -- exact elimination of slack variables from the reduced KKT system (SURVEY 8(f) rank 3,
// Lasso / Huber / portfolio-type problems: BASELINE config 3).
//
// A variable j is a SLACK when column j of A has exactly one entry a_j (in row r = r(j)), row and
// column j of P hold at most the diagonal entry P_jj, and no other slack shares row r.  Then the
// block K_YY of  K = P + sigma I + A' diag(rho) A  over the slack set Y is DIAGONAL,
//     D_j = P_jj + sigma + a_j^2 rho_r,
// every coupling of j goes through row r, and the Schur complement of K over the other variables X
// is the same kind of operator with a modified rho on the partner rows and the slack columns removed:
//     S = K_XX - K_XY D^-1 K_YX = P_XX + sigma I + A_X' diag(rho~) A_X,
//     rho~_r = rho_r (P_jj + sigma) / D_j   on partner rows,   rho~ = rho elsewhere.
// Lasso (reference docs/examples/lasso.rst: y = A x - b as equality rows with rho x 1e3 and y in the
// objective) has rho_r a_j^2 ~ 1e2..1e3 against P_jj = 1: K is ill-conditioned through exactly these
// rows, S is not (rho~ <= P_jj / a_j^2) -- Jacobi-preconditioned CG needs 4 iterations on S where it
// needs 23-28 on K (tests/manual_lasso_schur_proto.py).
//
// The reference's QDLDL factorises the full KKT matrix exactly (lin_sys/direct/qdldl/
// qdldl_interface.c:341-376), so it never sees this conditioning; here the elimination is what lets the
// CG behave like it.  It costs NO extra pass over the matrices: with p_Y = 0 the existing kernels
// compute S p_X (rows X of [P | A'] [p ; rho~ .* (A p)]), and everything that involves Y is one
// element per slack, through its partner row:
//   before the rhs pass   r0_Yj = -(q_j + P_jj x_j + a_j v_r),  v_r += rho_r a_j r0_Yj / D_j
//                         (folds  -K_XY D^-1 r0_Y  into the pass that computes r0_X)
//   inside the CG         Minv_j = 0  =>  z_j = p_j = 0;  r_j is forced to 0 (pcg_update, EpiRhs)
//   after the CG          d_Yj = (r0_Yj - a_j rho_r (A_X d_X)_r) / D_j,  x~_j = x_j + d_Yj,
//                         (A x~)_r += a_j d_Yj      (k_admm_update, by the thread of row r)
#pragma once
#include "dev_mats.cuh"
#include <vector>

struct SlackView {          // by value into kernels
  int ny;
  const int *slot;          // m: index into the lists below, -1 = row without a slack
  const int *var;           // ny: the slack variable of the row
  const double *a, *D;      // ny: its entry of A (scaled), D_j
  const double *r0;         // ny: r0_Yj of the current ADMM iteration
  const double *rho_cg;     // m: rho~
  const double *minv;       // n: the CG's inverse diagonal, 0 on the slacks
};

struct SlackDev {
  bool on = false;
  int ny = 0;
  int *slot = nullptr;                      // m
  int *row = nullptr, *var = nullptr;       // ny
  int *posA = nullptr, *posP = nullptr;     // ny: slots of a_j in A.val, of P_jj in M.val (-1: none)
  double *a = nullptr, *pjj = nullptr, *D = nullptr, *r0 = nullptr;   // ny
  double *rho_cg = nullptr;                 // m
  double *Minv_cg = nullptr;                // n
};

static inline SlackView slack_view(const SlackDev &s) {
  return SlackView{s.on ? s.ny : 0, s.slot, s.var, s.a, s.D, s.r0, s.rho_cg, s.Minv_cg};
}

// positions of a_j (row r of CSR(A)) and P_jj (P part of row j of CSR([P | A']))
__global__ void k_slack_locate(int ny, const int *row, const int *var, const int *Arp, const int *Acol,
                               const int *Mrp, const int *Mcol, const int *Msplit, int *posA, int *posP) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ny) return;
  const int r = row[t], j = var[t];
  int pa = -1, pp = -1;
  for (int e = Arp[r]; e < Arp[r + 1]; ++e) if (Acol[e] == j) { pa = e; break; }
  for (int e = Mrp[j]; e < Msplit[j]; ++e) if (Mcol[e] == j) { pp = e; break; }
  posA[t] = pa; posP[t] = pp;
}

// values that depend on rho / sigma / the (scaled) matrix entries; rho_cg holds a copy of rho on entry
__global__ void k_slack_refresh(int ny, const int *row, const int *posA, const int *posP, const double *Aval,
                                const double *Mval, const double *rho, double sigma, double *a, double *pjj,
                                double *D, double *rho_cg) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ny) return;
  const int r = row[t];
  const double aj = posA[t] >= 0 ? Aval[posA[t]] : 0.0;
  const double pj = posP[t] >= 0 ? Mval[posP[t]] : 0.0;
  const double rr = rho[r];
  const double d = pj + sigma + aj * aj * rr;
  a[t] = aj; pjj[t] = pj; D[t] = d;
  rho_cg[r] = rr * ((pj + sigma) / d);
}

__global__ void k_slack_mask(int ny, const int *var, double *Minv_cg) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < ny) Minv_cg[var[t]] = 0.0;
}

// before the rhs pass of an ADMM iteration (v = y + rho .* (A x - z) freshly written by its producer)
__global__ void k_slack_pre(int ny, const int *row, const int *var, const double *a, const double *pjj,
                            const double *D, const double *rho, const double *q, const double *x, double *v,
                            double *r0) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ny) return;
  const int r = row[t], j = var[t];
  const double aj = a[t], vr = v[r];
  const double ry = -(q[j] + pjj[t] * x[j] + aj * vr);
  r0[t] = ry;
  v[r] = vr + rho[r] * aj * ry / D[t];
}

static void slack_free(SlackDev &s) {
  dev_free(s.slot); dev_free(s.row); dev_free(s.var); dev_free(s.posA); dev_free(s.posP);
  dev_free(s.a); dev_free(s.pjj); dev_free(s.D); dev_free(s.r0); dev_free(s.rho_cg); dev_free(s.Minv_cg);
  s = SlackDev();
}

// Host: find the slack set from the caller's CSC arrays (P upper triangular).  Active when at least
// n/8 variables qualify (OSQP_B200_SLACK=0 disables, =1 lowers the bar to one variable).  0 on success.
static int slack_setup(SlackDev &s, const csc *P, const csc *A, const DevMats &W, cudaStream_t st) {
  const char *env = getenv("OSQP_B200_SLACK");
  if (env && env[0] == '0') return 0;
  const int n = W.n, m = W.m;
  if (!m || !n) return 0;
  std::vector<char> offd((size_t)n, 0);
  for (int c = 0; c < n; ++c)
    for (c_int e = P->p[c]; e < P->p[c + 1]; ++e) {
      const int i = (int)P->i[e];
      if (i != c) { offd[i] = 1; offd[c] = 1; }
    }
  std::vector<int> slot((size_t)m, -1), row, var;
  for (int j = 0; j < n; ++j) {
    if (offd[j] || A->p[j + 1] - A->p[j] != 1) continue;
    const int r = (int)A->i[A->p[j]];
    if (slot[r] >= 0) continue;                 // one slack per row
    slot[r] = (int)row.size();
    row.push_back(r); var.push_back(j);
  }
  const int ny = (int)row.size();
  const int bar = (env && env[0] == '1') ? 1 : (n + 7) / 8;
  if (ny < bar || ny == 0) return 0;
  s.ny = ny;
  if (dev_alloc(&s.slot, (size_t)m) || dev_alloc(&s.row, (size_t)ny) || dev_alloc(&s.var, (size_t)ny) ||
      dev_alloc(&s.posA, (size_t)ny) || dev_alloc(&s.posP, (size_t)ny) || dev_alloc(&s.a, (size_t)ny) ||
      dev_alloc(&s.pjj, (size_t)ny) || dev_alloc(&s.D, (size_t)ny) || dev_alloc(&s.r0, (size_t)ny) ||
      dev_alloc(&s.rho_cg, (size_t)m) || dev_alloc(&s.Minv_cg, (size_t)n)) { slack_free(s); return 1; }
  // pageable sources: the copies are complete (staged) when the calls return
  cudaMemcpyAsync(s.slot, slot.data(), sizeof(int) * m, cudaMemcpyHostToDevice, st);
  cudaMemcpyAsync(s.row, row.data(), sizeof(int) * ny, cudaMemcpyHostToDevice, st);
  cudaMemcpyAsync(s.var, var.data(), sizeof(int) * ny, cudaMemcpyHostToDevice, st);
  cudaMemsetAsync(s.r0, 0, sizeof(double) * ny, st);
  k_slack_locate<<<grid_for(ny), OSQP_CUDA_THREADS, 0, st>>>(ny, s.row, s.var, W.A.rowptr, W.A.col,
                                                             W.M.rowptr, W.M.col, W.M.split, s.posA, s.posP);
  if (cudaStreamSynchronize(st) != cudaSuccess || cudaGetLastError() != cudaSuccess) { slack_free(s); return 1; }
  s.on = true;
  return 0;
}
