#!/bin/bash
# usage: scripts/build_variant.sh NAME "-DXG_GEAR_WARPS=6 ..."   ->  variants_tmp/NAME/libxetgpu.so
# Experiment builds of the library with different kernel geometries (loaded through XETGPU_LIBRARY).
set -e
cd "$(dirname "$0")/.."
name=$1; shift
out=variants_tmp/$name
mkdir -p $out
NVCC=/usr/local/cuda/bin/nvcc
FLAGS="-O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-Wall,-Wno-unused-function --expt-relaxed-constexpr"
objs=""
for f in gear resolve chunk_hash merkle merkle_static safetensors api; do
  $NVCC $FLAGS "$@" -c ome_b200/csrc/$f.cu -o $out/$f.o 2> $out/$f.log &
  objs="$objs $out/$f.o"
done
wait
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o $out/libxetgpu.so $objs -lcudart_static -lpthread -ldl -lrt
echo "$name: $*" > $out/FLAGS
