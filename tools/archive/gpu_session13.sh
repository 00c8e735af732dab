#!/bin/bash
mkdir -p gpurun_out
export GEOTX_LIB=$PWD/tools/bin/abl_noprod.so K2_SCORES=1
timeout 120 python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s13_plain.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on -k regex:gse_embed_cheb -s 1 -c 1 -f -o gpurun_out/r02_k3r_noprod python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s13_ncu.log 2>&1
echo rc=$?; tail -3 gpurun_out/s13_ncu.log
