#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_patch_scores.py tests/test_tail_chain.py -m gpu -q --timeout 300 > gpurun_out/test_patch_scores.log 2>&1
echo "test_patch_scores exit $?" | tee -a gpurun_out/summary.txt
tail -n 30 gpurun_out/test_patch_scores.log
timeout 300 python - > gpurun_out/k12_timing.json 2> gpurun_out/k12_timing.err <<'PY'
import json, sys
sys.argv = ['bench.py']
import bench, pkgpath
pkg = pkgpath.load_package()
print(json.dumps({s: bench.time_patch_scores(pkg, s) for s in ('3dmatch', 'kitti')}, indent=1))
PY
echo "k12_timing exit $?"; cat gpurun_out/k12_timing.json; tail -n 5 gpurun_out/k12_timing.err
timeout 300 python tools/tail_bench.py > gpurun_out/tail_bench.json 2> gpurun_out/tail_bench.err
echo "tail_bench exit $?"; cat gpurun_out/tail_bench.json; tail -n 5 gpurun_out/tail_bench.err
