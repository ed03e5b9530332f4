#!/bin/bash
# builds the SELL kernel variants of spmv_sell_bench.cu into bin/ (git-ignored, travels to the GPU box)
cd "$(dirname "$0")"; mkdir -p bin; rm -f bin/sell_*
build() { # tma ch nst minb lgm
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -I../../include \
    -DSELL_TMA=$1 -DSELL_CH=$2 -DSELL_NST=$3 -DSELL_MINB=$4 -DSELL_LGM=$5 -DSELL_PF=${6:-1} -DSELL_STREAM_LD=${7:-0} spmv_sell_bench.cu \
    -o bin/sell_t$1_c$2_s$3_b$4_g$5_p${6:-1}_l${7:-0} &
}
build 0 4 1 4 2 1 0; build 0 4 1 4 2 1 1; build 0 5 1 4 2 1 1; build 0 8 1 3 2 1 1
build 0 4 1 4 3 1 0; build 0 4 1 4 4 1 0; build 0 4 1 4 5 1 0; build 0 4 1 4 1 1 0
wait; ls bin
