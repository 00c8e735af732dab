#!/bin/bash
mkdir -p gpurun_out
timeout 180 python -m pytest tests/test_gpu_k2_embed.py -m gpu -q -x 2>&1 | tail -12 > gpurun_out/s7_tests_k2.txt; echo "k2 tests rc=$?"; tail -5 gpurun_out/s7_tests_k2.txt
K2_QUICK=1 K2_OUT=k2q_v6.json timeout 180 python tools/k2_embed_bench.py > gpurun_out/k2q_v6.log 2>&1; echo "bench rc=$?"; tail -3 gpurun_out/k2q_v6.log | cut -c1-420
timeout 300 python -m pytest tests -m gpu -q 2>&1 | tail -12 > gpurun_out/s7_tests.txt; tail -5 gpurun_out/s7_tests.txt
K2_SCORES=1 timeout 120 python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s7_plain1.log 2>&1 && K2_SCORES=1 timeout 300 ncu --set full --clock-control none --import-source on -k regex:gse_embed_cheb -s 1 -c 1 -f -o gpurun_out/r02_k3r_v6 python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s7_ncu1.log 2>&1
timeout 400 python bench.py --steps 20 --warmup 5 --no-sweep > gpurun_out/r02_bench_v5.json 2> gpurun_out/r02_bench_v5.err; echo "bench rc=$?"; tail -c 400 gpurun_out/r02_bench_v5.err; head -c 300 gpurun_out/r02_bench_v5.json
