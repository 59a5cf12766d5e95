#!/bin/bash
# Builds a tuning variant of libvfcuda.so into build/variants/NAME/ (git-ignored, travels with gpurun);
# run it with LD_LIBRARY_PATH=build/variants/NAME (the host module finds the library through RUNPATH,
# which LD_LIBRARY_PATH precedes).  usage: tools/build_variant.sh NAME "-DVF_GM_CTAS=3 -DVF_GM_ROWS=15"
set -e
NAME=$1; shift
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/build/variants/$NAME
mkdir -p $OUT
C=$ROOT/vulkan-forge_b200/csrc
/usr/local/cuda/bin/nvcc $@ -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 --fmad=false -std=c++17 -ccbin /usr/bin/g++ \
  -Xcompiler -fPIC,-ffp-contract=off -shared -o $OUT/libvfcuda.so $C/vf_context.cu $C/vf_terrain.cu $C/vf_dem.cu $C/vf_misc.cu $C/vf_png_dev.cu $C/vf_png.cpp -lz
echo built $OUT/libvfcuda.so
