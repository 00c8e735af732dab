#!/bin/bash
mkdir -p gpurun_out
TAG=${TAG:-v10}
summ() { python - "$1" <<'PY'
import json,sys
r=json.load(open(sys.argv[1]))
for k,v in r.items(): print(k,'E-store %.4f | K3r %.4f | K3r split %.4f ms (%.1f clk/pair)'%(v['cheb']['ms'],v['cheb_scores_K3r']['ms'],v['cheb_scores_K3r_split']['ms'],v['cheb_scores_K3r_split']['clk_per_pair_per_sm_at_1965MHz']))
PY
}
timeout 300 python -m pytest tests/test_gpu_k2_embed.py -m gpu -q -x 2>&1 | tail -3
K2_QUICK=1 K2_OUT=k2q_$TAG.json timeout 180 python tools/k2_embed_bench.py > gpurun_out/k2q_$TAG.log 2>&1; echo "== $TAG rc=$?"; summ gpurun_out/k2q_$TAG.json
for v in $VARIANTS; do
GEOTX_LIB=$PWD/tools/bin/$v.so K2_QUICK=1 K2_OUT=k2q_$v.json timeout 180 python tools/k2_embed_bench.py > gpurun_out/k2q_$v.log 2>&1; echo "== $v rc=$?"; summ gpurun_out/k2q_$v.json
done
timeout 400 python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/s12_tests.txt; tail -3 gpurun_out/s12_tests.txt
timeout 600 python bench.py --no-sweep > gpurun_out/r02_bench_$TAG.json 2> gpurun_out/r02_bench_$TAG.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02_bench_$TAG.err; head -c 200 gpurun_out/r02_bench_$TAG.json; echo
if [ -n "$AB" ]; then
GEOTX_PARALLEL_CLOUDS=0 timeout 600 python bench.py --no-sweep --no-cpu-baseline > gpurun_out/r02_bench_${TAG}_serial.json 2> gpurun_out/r02_bench_${TAG}_serial.err; echo "bench (one stream) rc=$?"; head -c 200 gpurun_out/r02_bench_${TAG}_serial.json; echo
fi
