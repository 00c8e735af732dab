#!/bin/bash
mkdir -p gpurun_out
timeout 200 python bench.py --launch-list --steps 2 > gpurun_out/s17_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/r02_forward_full python bench.py --launch-list --steps 1 > gpurun_out/s17_ncu.log 2>&1
echo rc=$?; tail -2 gpurun_out/s17_ncu.log; ls -la gpurun_out/r02_forward_full.ncu-rep
