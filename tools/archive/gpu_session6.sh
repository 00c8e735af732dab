#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_k2_embed.py tests/test_gpu_forward.py -m gpu -q 2>&1 | tail -12 > gpurun_out/s6_tests.txt; tail -5 gpurun_out/s6_tests.txt
K2_QUICK=1 K2_OUT=k2q_v5.json python tools/k2_embed_bench.py > gpurun_out/k2q_v5.log 2>&1; echo "== base (3 CTAs, unroll 2)"; tail -3 gpurun_out/k2q_v5.log | cut -c1-420
for v in cheb_c4 cheb_c2 cheb_c3u1 cheb_c3u4; do
  GEOTX_LIB=$PWD/tools/bin/$v.so K2_QUICK=1 K2_OUT=k2q_$v.json python tools/k2_embed_bench.py > gpurun_out/k2q_$v.log 2>&1
  echo "== $v"; tail -3 gpurun_out/k2q_$v.log | cut -c1-420
done
K2_SCORES=1 python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s6_plain1.log 2>&1 && K2_SCORES=1 ncu --set full --clock-control none --import-source on -k regex:gse_embed_cheb -s 1 -c 1 -f -o gpurun_out/r02_k3r_v5 python tools/k2_embed_bench.py once cheb 1024 256 > gpurun_out/s6_ncu1.log 2>&1
python bench.py --steps 20 --warmup 5 --no-sweep > gpurun_out/r02_bench_v4.json 2> gpurun_out/r02_bench_v4.err; echo "bench rc=$?"; tail -c 500 gpurun_out/r02_bench_v4.err; head -c 300 gpurun_out/r02_bench_v4.json
