#!/bin/bash
for v in "" k13_fence k13_stay k13_one; do
  echo "== variant '$v'"
  if [ -n "$v" ]; then export GEOTX_LIB=$PWD/tools/bin/$v.so; fi
  timeout 200 python tools/dense_stress.py 2>&1 | grep -v "iter" | head -8
done
