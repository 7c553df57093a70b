#!/bin/bash
# usage: tools/build_variant.sh <name> [-DFLAG ...]  -> mobileraytracer_b200/libmobilert_b200_<name>.so (A/B experiments; select with MRT_B200_LIB)
N=$1; shift
cd "$(dirname "$0")/../mobileraytracer_b200/csrc" && nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC,-O2 \
  -ccbin /usr/bin/g++ "$@" -shared -o ../libmobilert_b200_$N.so mrt_bvh_build.cu mrt_render.cu mrt_wavefront.cu mrt_api.cu 2>&1 | grep -i " error"
