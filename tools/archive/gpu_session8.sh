#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_k2_embed.py -m gpu -q -x -s 2>&1 | grep -E "split|passed|failed|Error|error|assert" | tail -14; 
K2_QUICK=1 K2_OUT=k2q_v7.json timeout 180 python tools/k2_embed_bench.py > gpurun_out/k2q_v7.log 2>&1; echo "bench rc=$?"; python - <<'PY'
import json
r=json.load(open('gpurun_out/k2q_v7.json'))
for k,v in r.items(): print(k,'E-store %.4f | K3r %.4f | K3r split %.4f ms (%.1f clk/pair, %.0f TF)'%(v['cheb']['ms'],v['cheb_scores_K3r']['ms'],v['cheb_scores_K3r_split']['ms'],v['cheb_scores_K3r_split']['clk_per_pair_per_sm_at_1965MHz'],v['cheb_scores_K3r_split']['executed_tflops']))
PY
timeout 300 python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/s8_tests.txt; tail -4 gpurun_out/s8_tests.txt
K2_SCORES=1 timeout 120 python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s8_plain1.log 2>&1 && K2_SCORES=1 timeout 300 ncu --set full --clock-control none --import-source on -k regex:gse_embed_cheb -s 1 -c 1 -f -o gpurun_out/r02_k3r_v7 python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s8_ncu1.log 2>&1
timeout 400 python bench.py --steps 20 --warmup 5 --no-sweep > gpurun_out/r02_bench_v6.json 2> gpurun_out/r02_bench_v6.err; echo "bench rc=$?"; tail -c 400 gpurun_out/r02_bench_v6.err; head -c 300 gpurun_out/r02_bench_v6.json
