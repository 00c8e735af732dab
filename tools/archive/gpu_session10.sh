#!/bin/bash
mkdir -p gpurun_out
summ() { python - "$1" <<'PY'
import json,sys
r=json.load(open(sys.argv[1]))
for k,v in r.items(): print(k,'E-store %.4f | K3r %.4f | K3r split %.4f ms (%.1f clk/pair, %.0f TF)'%(v['cheb']['ms'],v['cheb_scores_K3r']['ms'],v['cheb_scores_K3r_split']['ms'],v['cheb_scores_K3r_split']['clk_per_pair_per_sm_at_1965MHz'],v['cheb_scores_K3r_split']['executed_tflops']))
PY
}
K2_QUICK=1 K2_OUT=k2q_v9.json timeout 180 python tools/k2_embed_bench.py > gpurun_out/k2q_v9.log 2>&1; echo "== distance HMMAs hoisted rc=$?"; summ gpurun_out/k2q_v9.json
timeout 400 python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/s10_tests.txt; tail -3 gpurun_out/s10_tests.txt
timeout 900 python bench.py > gpurun_out/r02_bench_v9.json 2> gpurun_out/r02_bench_v9.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02_bench_v9.err; head -c 200 gpurun_out/r02_bench_v9.json; echo
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_ref.json 2> gpurun_out/r02_bench_ref.err; echo "ref rc=$?"; head -c 600 gpurun_out/r02_bench_ref.json; echo
timeout 200 python bench.py --launch-list --steps 2 > gpurun_out/s10_plain_ll.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02_launches_3dmatch_1024.csv python bench.py --launch-list --steps 2 > gpurun_out/s10_ncu_ll.log 2>&1; echo "launch list rc=$?"; wc -l gpurun_out/r02_launches_3dmatch_1024.csv
