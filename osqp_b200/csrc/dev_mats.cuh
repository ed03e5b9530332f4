// dev_mats.cuh -- device-resident problem matrices.
//
// Layout in HBM (int32 indices, fp64 values, 12 B per stored nonzero):
//   A  : CSR, m x n               -- rows of A, for A*x
//   M  : CSR, n x (n+m) = [Pf | A'] -- row j = row j of the FULL symmetric P followed
//        by column j of A (the caller's CSC of A re-read as CSR of A'), for
//        P*x + A'*y in ONE pass; columns >= n gather from the m-vector.
//   Araw / Praw : the caller's CSC value arrays, un-scaled, kept so that
//        osqp_update_P/A and re-equilibration restart from the original data
//        (reference src/osqp.c:1012-1279 un-scales and re-scales in place instead).
// The reference's counterparts are csc P, A (include/types.h:21-29) and the
// KKT index maps of src/kkt.c:6-177; here the maps are CsrDev::src.
#pragma once
#include "spmv_direct.cuh"
#include "spmv_sell.cuh"
#include "spmv_tile.cuh"

struct DevMats {
  int n = 0, m = 0, nnzA = 0, nnzP = 0, nnzPf = 0;
  CsrDev A;                 // m x n
  CsrDev M;                 // n x (n+m)
  double *Araw = nullptr;   // nnzA, CSC order
  double *Praw = nullptr;   // nnzP, CSC order (upper triangle)
  int refcount = 1;
};

// Build everything from host CSC inputs (c_int = 64-bit indices).  0 on success.
int mats_build(DevMats &W, const csc *P, const csc *A, cudaStream_t st);
void mats_free(DevMats &W);
// Overwrite raw values from HOST arrays: all of them (idx == NULL) or by index list.
int mats_update_raw(DevMats &W, bool isP, const c_float *vals, const c_int *idx, c_int count,
                    cudaStream_t st);
// A.val / M.val <- raw (un-scaled) values.
int mats_load_raw(DevMats &W, cudaStream_t st);
// out[row] = max |val| over the row (out must be zeroed).  P-only variant for M.
void mats_row_absmax(const CsrDev &C, double *out, bool p_part_only, int n, cudaStream_t st);
// A <- diag(e) A diag(d);  M: P part <- diag(d) P diag(d), A' part <- diag(d) A' diag(e)
void mats_scale(DevMats &W, const double *d, const double *e, cudaStream_t st);
// P part of M *= *c_dev
void mats_scale_P(DevMats &W, const double *c_dev, cudaStream_t st);
// Re-read the values of the sliced-ELL mirrors (spmv_sell.cuh) from the CSR arrays if a
// scaling pass / matrix update changed them; until then SpMV launches use the CSR kernels.
void mats_sell_sync(DevMats &W, cudaStream_t st);
