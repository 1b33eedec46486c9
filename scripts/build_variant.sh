#!/bin/bash
# Build a tuning variant of the library:  scripts/build_variant.sh <tag> "<extra nvcc flags>"   ->  ray-tracer_b200/build/variants/librt_b200_<tag>.so
# (select it with RT_B200_LIB=<path>; the default library is never touched)
set -e
cd "$(dirname "$0")/../ray-tracer_b200"
TAG=$1; FLAGS=$2
mkdir -p build/variants/$TAG
for f in kernels capi lbvh; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC $FLAGS -c csrc/$f.cu -o build/variants/$TAG/$f.o &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/variants/librt_b200_$TAG.so build/variants/$TAG/kernels.o build/variants/$TAG/capi.o build/variants/$TAG/lbvh.o build/bvh_build.o build/host_capi.o -cudart static -lz -ldl
echo built build/variants/librt_b200_$TAG.so
