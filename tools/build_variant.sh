#!/bin/bash
# Build a variant of libshk_b200.so with extra nvcc flags for A/B measurements:
#   tools/build_variant.sh NAME "-DSHK_SPLIT_MASK=15u"  ->  gpurun_variants/libshk_NAME.so   (use with SHK_LIB=...)
set -e
cd "$(dirname "$0")/../sparrowhawk_b200/csrc"
name=$1; shift
out=../../gpurun_variants; mkdir -p $out/obj_$name
ARCH="-gencode arch=compute_100a,code=sm_100a"
FL="$ARCH -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-Wall,-Wno-unused-function $*"
for f in shk_kernels.cu shk_api.cu shk_shard.cu shk_noalloc.cu shk_host.cpp shk_clean.cpp; do nvcc $FL -c $f -o $out/obj_$name/${f%.*}.o & done; wait
nvcc $ARCH -shared -o $out/libshk_$name.so $out/obj_$name/*.o -lz -cudart static
rm -rf $out/obj_$name
echo built $out/libshk_$name.so
