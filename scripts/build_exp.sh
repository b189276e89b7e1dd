#!/bin/bash
# Fast experimental build of the product library (two kernel variants only) into sspp_b200/csrc/libexp_<tag>.so,
# plus the replica oracle.  Use with SSPP_CUDA_LIB=<that file>.   usage: bash scripts/build_exp.sh <tag> [extra nvcc flags]
set -e
TAG=$1; shift
cd "$(dirname "$0")/../sspp_b200/csrc"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 --fmad=false -std=c++17 -Xcompiler -fPIC,-fno-fast-math,-ffp-contract=off -DSSPP_FAST_BUILD "$@" \
  -I ../../include -shared -Xlinker -Bsymbolic -o libexp_$TAG.so sspp_cabi.cu sspp_kernels.cu sspp_model.cpp sspp_spline.cpp
make -s -C ../../oracle
echo built libexp_$TAG.so
