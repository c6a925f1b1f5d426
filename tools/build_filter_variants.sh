#!/bin/bash
# Build librt_b200 variants of the filter kernel configuration into build/ (A/B on the GPU box via RT_B200_LIB).
# usage: tools/build_filter_variants.sh "RPT OPI MINB TILE" ...
cd "$(dirname "$0")/../opencl-raytracer_b200" || exit 1
mkdir -p ../build
for cfg in "$@"; do
  set -- $cfg
  out=../build/librt_b200_f$1_$2_$3_$4.so
  nvcc -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-Wall,-Wno-unused-function --expt-relaxed-constexpr \
    -DRT_FILTER_RPT=$1 -DRT_FILTER_OPI=$2 -DRT_FILTER_MINB=$3 -DRT_FILTER_TILE=$4 -shared -o $out \
    csrc/rt_api.cu csrc/rt_kernels.cu csrc/rt_wavefront.cu csrc/rt_lbvh.cu csrc/rt_scene.cpp 2>&1 | grep -i "error" 
  echo "built $out"
done
