#!/bin/bash
# Compiles the reference's own CPU collate operators (torch-free C++: three source files + the vendored nanoflann
# header) from /root/reference into oracle/_ref/libgeotx_ref_cpu.so, behind the plain-C shim oracle/ref_cpu_shim.cpp.
# No reference source is copied; g++ reads the files where they lie.  -ffp-contract=off: the reference's own build
# (setup.py, torch cpp_extension defaults, x86-64 baseline) has no FMA either.
set -e
REF=${GEOTX_REFERENCE_ROOT:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
EXT=$REF/geotransformer/extensions
mkdir -p "$HERE/_ref"
g++ -O2 -std=c++14 -ffp-contract=off -fPIC -shared -I"$REF" \
  "$HERE/ref_cpu_shim.cpp" \
  "$EXT/cpu/radius_neighbors/radius_neighbors_cpu.cpp" \
  "$EXT/cpu/grid_subsampling/grid_subsampling_cpu.cpp" \
  "$EXT/extra/cloud/cloud.cpp" \
  -o "$HERE/_ref/libgeotx_ref_cpu.so"
echo "$HERE/_ref/libgeotx_ref_cpu.so"
