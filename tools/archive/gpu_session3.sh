#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -25 > gpurun_out/s3_tests.txt; tail -8 gpurun_out/s3_tests.txt
K2_SCORES=1 python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s3_plain1.log 2>&1 && K2_SCORES=1 ncu --set full --clock-control none --import-source on -k regex:gse_embed_cheb -s 1 -c 1 -f -o gpurun_out/r02_k3r python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s3_ncu1.log 2>&1
python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s3_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gse_embed_cheb -s 1 -c 1 -f -o gpurun_out/r02_k2c python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s3_ncu2.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -3 gpurun_out/s3_ncu1.log
