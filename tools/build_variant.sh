#!/bin/bash
# Build a variant of libdibr_b200.so with extra -D flags into build/variants/<name>.so (run A/B on the GPU box through
# DIBR_B200_LIB=$PWD/build/variants/<name>.so).   bash tools/build_variant.sh bwd4 -DDIBR_RASTER_BWD_CTAS=4
N=$1; shift
mkdir -p build/variants
nvcc -shared -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -cudart shared \
  -Xcompiler -fPIC,-fvisibility=hidden -ldl "$@" -o build/variants/$N.so kaolin_b200/csrc/dibr_c_abi.cu && echo "built build/variants/$N.so"
