#!/bin/bash
# GPU box: K2 time of every variant library under tools/bin/variants (first case of tools/k2_bench.py only)
for f in tools/bin/variants/lib_*.so; do
  echo -n "$(basename $f .so | sed 's/^lib_//'): "
  K2_CASE=${K2_CASE:-0} timeout 120 python tools/k2_bench.py $f 2>&1 | tail -1
done
