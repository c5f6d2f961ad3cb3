#!/bin/bash
# usage: build_harness_variant.sh <suffix> <extra nvcc flags...>   -> build/test_igemm2_<suffix>
set -e
cd "$(dirname "$0")/.."
SUF=$1; shift
FL="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17"
mkdir -p build/var_$SUF
for f in igemm_sm100 igemm2_sm100 runtime; do nvcc $FL "$@" -c -o build/var_$SUF/$f.o csrc/$f.cu & done; wait
nvcc $FL "$@" -o build/test_igemm2_$SUF csrc/test_igemm2.cu build/var_$SUF/*.o -lcuda
echo build/test_igemm2_$SUF
