#!/bin/bash
# Experiment helper: build plugin variants with different compile-time tuning into gpurun_out-independent dirs
# usage: tools/build_variants.sh name "EXTRA flags" [name "flags" ...]
cd "$(dirname "$0")/../osqp_b200/csrc"
while [ $# -gt 1 ]; do
  name=$1; flags=$2; shift 2
  ( make -s LIB=../lib_$name EXTRA="$flags" ../lib_$name/libosqp_cuda_pcg.so > /dev/null 2>&1; echo "$name done rc=$?" ) &
done
wait
