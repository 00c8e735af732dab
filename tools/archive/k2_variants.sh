#!/bin/bash
# Builds timing-experiment variants of the library (K2_DBG_* switches; wrong results by design) into tools/bin/variants.
set -e
cd "$(dirname "$0")/.."
mkdir -p tools/bin/variants
SRC=ref-geotransformer_b200/csrc
for v in "$@"; do
  flags=""
  # NAME -> -DK2_DBG_NAME ; NAME=VALUE -> -DK2_NAME=VALUE (e.g. PGROUPS=2)
  for f in $(echo $v | tr '+' ' '); do
    case $f in *=*) flags="$flags -DK2_$f";; *) flags="$flags -DK2_DBG_$f";; esac
  done
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Xcompiler -fPIC -shared $flags \
    $SRC/*.cu \
    -o tools/bin/variants/lib_$v.so -lcudart &
done
wait
ls tools/bin/variants
