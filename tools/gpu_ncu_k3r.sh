#!/bin/bash
# one --set full capture (with SASS-level sampling) of K3r alone: split instance, n = 1024, C = 256
mkdir -p gpurun_out
export K2_SCORES=1
timeout 120 python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/k3r_plain.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on -k regex:gse_embed_cheb -s 1 -c 1 -f -o gpurun_out/r02_k3r_final python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/k3r_ncu.log 2>&1
echo rc=$?; tail -2 gpurun_out/k3r_ncu.log
