#!/bin/bash
# usage: tools/build_variant.sh name [-DRM_...=...]...   -> build/variants/lib_<name>.so (sweeps of launch-shape macros)
name=$1; shift
cd "$(dirname "$0")/../ray_marching_b200/csrc" && mkdir -p ../../build/variants && \
nvcc -shared -o ../../build/variants/lib_$name.so -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 \
  -Xcompiler -fPIC --expt-relaxed-constexpr "$@" rm_host.cu rm_forward.cu rm_backward.cu rm_probe.cu
