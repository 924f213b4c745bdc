#!/bin/bash
# Builds an experiment variant of the library: scripts/build_variant.sh <name> <extra nvcc flags...>  -> build/variants/lib<name>.so
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p build/variants
C=ignis-masterthesis_b200/csrc
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC -shared "$@" \
  -o build/variants/lib$name.so $C/bvh8_build.cpp $C/debug_api.cpp $C/trace.cu $C/wavefront.cu $C/tonemap.cu $C/device.cu $C/multi_gpu.cpp -ldl 2>&1 | grep -i "error" || true
ls -la build/variants/lib$name.so
