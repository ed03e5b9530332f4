#!/bin/bash
# Experiment: time the SpMV kernels of tuning variants built by tools/build_variants.sh
# (run on the GPU box from the repo root)
run() {  # name lib cps gmul
  echo "=== $1 lib=$2 cps=$3 gmul=$4"
  OSQP_CUDA_PCG_LIB=$PWD/osqp_b200/$2/libosqp_cuda_pcg.so OSQP_B200_CPS=$3 OSQP_B200_GMUL=$4 \
    timeout 120 python tools/spmv_phases.py 0 2>&1 | grep -E "spmv_A|spmv_M|rhs|admm_update|resid"
}
run base lib 4 1
run base_g2 lib 4 2
run base_cps8 lib 8 1
run tb8m3 lib_tb8m3 3 1
run tb8m3_g2 lib_tb8m3 3 2
run tb8m4 lib_tb8m4 4 1
run tb4m5 lib_tb4m5 5 1
run tb4m5_g2 lib_tb4m5 5 2
run tb6m4 lib_tb6m4 4 1
run tb2m6 lib_tb2m6 6 1
run tb2m6_g2 lib_tb2m6 6 2
