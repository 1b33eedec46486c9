#!/bin/bash
# Build tuning variants of librt_b200.so into ray-tracer_b200/build/variants/<name>.so:  scripts/build_variants.sh name "-DFLAG=..." [name flags]...
set -e
cd "$(dirname "$0")/../ray-tracer_b200"
mkdir -p build/variants
make -s librt_b200.so > /dev/null
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC $flags -c csrc/kernels.cu -o build/variants/kernels_$name.o
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/variants/$name.so build/variants/kernels_$name.o build/capi.o build/lbvh.o build/bvh_build.o build/host_capi.o -cudart static -lz
  echo built $name "($flags)"
done
