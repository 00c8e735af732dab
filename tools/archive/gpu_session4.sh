#!/bin/bash
mkdir -p gpurun_out
python tools/debug_modes.py > gpurun_out/s4_debug.txt 2>&1; tail -30 gpurun_out/s4_debug.txt
python -m pytest tests -m gpu -q 2>&1 | tail -25 > gpurun_out/s4_tests.txt; tail -8 gpurun_out/s4_tests.txt
K2_QUICK=1 K2_OUT=k2q_v3.json python tools/k2_embed_bench.py > gpurun_out/k2q_v3.log 2>&1; tail -3 gpurun_out/k2q_v3.log | cut -c1-700
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_v2.json 2> gpurun_out/r02_bench_v2.err; echo "bench rc=$?"; tail -c 800 gpurun_out/r02_bench_v2.err; head -c 2500 gpurun_out/r02_bench_v2.json
