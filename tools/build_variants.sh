#!/bin/bash
# Builds timing-experiment variants of lib/libgeotx_b200.so into tools/bin/ (git-ignored; travels to the GPU box):
#   tools/build_variants.sh <name> <file.cu> "<-D flags>" [...]
# Each variant relinks the in-tree objects with ONE source recompiled under the given macros; run with GEOTX_LIB=tools/bin/<name>.so
set -e
cd "$(dirname "$0")/.."
mkdir -p tools/bin
PKG=ref-geotransformer_b200
while [ $# -ge 3 ]; do
  name=$1; src=$2; flags=$3; shift 3
  obj=tools/bin/${name}_$(basename "$src" .cu).o
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $flags -c $PKG/csrc/$src -o $obj
  others=$(ls $PKG/build/*.o | grep -v "/$(basename "$src" .cu).o")
  nvcc -shared -o tools/bin/$name.so $obj $others -gencode arch=compute_100a,code=sm_100a -lcudart
  echo "built tools/bin/$name.so ($flags)"
done
