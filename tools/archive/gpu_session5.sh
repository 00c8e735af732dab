#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -25 > gpurun_out/s5_tests.txt; tail -6 gpurun_out/s5_tests.txt
K2_QUICK=1 K2_OUT=k2q_v4.json python tools/k2_embed_bench.py > gpurun_out/k2q_v4.log 2>&1; tail -3 gpurun_out/k2q_v4.log | cut -c1-420
K2_SCORES=1 python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s5_plain1.log 2>&1 && K2_SCORES=1 ncu --set full --clock-control none --import-source on -k regex:gse_embed_cheb -s 1 -c 1 -f -o gpurun_out/r02_k3r_v4 python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s5_ncu1.log 2>&1
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_v3.json 2> gpurun_out/r02_bench_v3.err; echo "bench rc=$?"; tail -c 500 gpurun_out/r02_bench_v3.err; head -c 300 gpurun_out/r02_bench_v3.json
