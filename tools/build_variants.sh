#!/bin/bash
# Builds experiment variants of libpt_b200.so into build/ (git-ignored): launch shape x register cap.
# usage: tools/build_variants.sh "<threads> <min_blocks> [extra nvcc flags]" ...
set -e
cd "$(dirname "$0")/.."
mkdir -p build
for spec in "$@"; do
  set -- $spec
  t=$1; b=$2; shift 2
  tagx=$(echo "$*" | tr -c 'A-Za-z0-9\n' '_')
  out=build/libpt_b200_t${t}_b${b}${tagx:+_$tagx}.so
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -Iinclude \
    --expt-relaxed-constexpr -DPT_RENDER_THREADS=$t -DPT_RENDER_MIN_BLOCKS=$b "$@" -Xptxas -v -shared -o $out pathtracer_b200/csrc/pt_api.cu 2> $out.log
  echo "$out: $(grep -A2 'k_renderILj48' $out.log | grep -o 'Used [0-9]* registers') $(grep -A1 'k_renderILj48' $out.log | grep -o '[0-9]* bytes spill stores')"
done
