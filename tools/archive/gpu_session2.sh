#!/bin/bash
# GPU session: full -m gpu suite, K2c / K3r variants, bench line.  Outputs under gpurun_out/.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/s2_tests.txt; tail -6 gpurun_out/s2_tests.txt
K2_QUICK=1 K2_OUT=k2q_base.json python tools/k2_embed_bench.py > gpurun_out/k2q_base.log 2>&1; tail -3 gpurun_out/k2q_base.log | cut -c1-900
for v in cheb_u1 cheb_u4 cheb_fatan cheb_c1u4; do
  GEOTX_LIB=$PWD/tools/bin/$v.so K2_QUICK=1 K2_OUT=k2q_$v.json python tools/k2_embed_bench.py > gpurun_out/k2q_$v.log 2>&1
  echo "== $v"; tail -3 gpurun_out/k2q_$v.log | cut -c1-900
done
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_v1.json 2> gpurun_out/r02_bench_v1.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/r02_bench_v1.err; head -c 6000 gpurun_out/r02_bench_v1.json
