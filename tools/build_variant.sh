#!/bin/bash
# usage: tools/build_variant.sh NAME [-DMACRO=VALUE ...]  ->  nusiprop_b200/lib/variants/libnusiprop_b200_NAME.so
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p nusiprop_b200/lib/variants
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC,-ffp-contract=off -shared "$@" \
  -o nusiprop_b200/lib/variants/libnusiprop_b200_$name.so nusiprop_b200/csrc/nusi_capi.cu nusiprop_b200/csrc/nusi_host.cpp -Xptxas -v 2>&1 | grep -A2 "alpha_group\|alpha_compact" | grep "Used\|spill"
