#!/bin/bash
# builds the SELL kernel variants of spmv_sell_bench.cu into bin/ (git-ignored, travels to the GPU box)
cd "$(dirname "$0")"; mkdir -p bin; rm -f bin/sell_*
build() { # tma ch nst minb lgm
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -I../../include \
    -DSELL_TMA=$1 -DSELL_CH=$2 -DSELL_NST=$3 -DSELL_MINB=$4 -DSELL_LGM=$5 spmv_sell_bench.cu \
    -o bin/sell_t$1_c$2_s$3_b$4_g$5 &
}
build 1 2 3 4 2; build 1 2 4 4 2; build 1 1 4 4 2; build 1 1 6 4 2; build 1 2 3 3 2; build 1 4 3 3 2; build 1 3 3 4 2
build 0 8 1 4 2; build 0 6 1 4 2; build 0 10 1 4 2; build 0 5 1 4 2
wait; ls bin
