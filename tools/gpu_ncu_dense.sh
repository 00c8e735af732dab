#!/bin/bash
# --set full captures of the dense glue inside one eager forward (3DMatch 1024): K13 and K4f launches of the first blocks
mkdir -p gpurun_out
timeout 200 python bench.py --launch-list --steps 1 > gpurun_out/dense_plain.log 2>&1 && timeout 500 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"post_attention|linear_tile" -c 7 -f -o gpurun_out/r02_dense python bench.py --launch-list --steps 1 > gpurun_out/dense_ncu.log 2>&1
echo rc=$?; tail -2 gpurun_out/dense_ncu.log
