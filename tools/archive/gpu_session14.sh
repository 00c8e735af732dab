#!/bin/bash
mkdir -p gpurun_out
summ() { python - "$1" <<'PY'
import json,sys
r=json.load(open(sys.argv[1]))
for k,v in r.items(): print(k,'E-store %.4f | K3r %.4f | K3r split %.4f ms (%.1f clk/pair)'%(v['cheb']['ms'],v['cheb_scores_K3r']['ms'],v['cheb_scores_K3r_split']['ms'],v['cheb_scores_K3r_split']['clk_per_pair_per_sm_at_1965MHz']))
PY
}
for v in $VARIANTS; do
GEOTX_LIB=$PWD/tools/bin/$v.so K2_QUICK=1 K2_OUT=k2q_$v.json timeout 180 python tools/k2_embed_bench.py > gpurun_out/k2q_$v.log 2>&1; echo "== $v rc=$?"; summ gpurun_out/k2q_$v.json
done
