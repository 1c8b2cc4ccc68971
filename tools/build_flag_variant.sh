#!/bin/bash
# Experiment build of libpt_b200.so with extra nvcc flags only (launch shape and register caps as in the product build).
# usage: tools/build_flag_variant.sh <tag> [nvcc flags...]   ->  build/libpt_b200_<tag>.so
set -e
cd "$(dirname "$0")/.."
mkdir -p build
tag=$1; shift
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -Iinclude \
  --expt-relaxed-constexpr "$@" -Xptxas -v -shared -o build/libpt_b200_$tag.so pathtracer_b200/csrc/pt_api.cu 2> build/libpt_b200_$tag.log
echo "build/libpt_b200_$tag.so"
