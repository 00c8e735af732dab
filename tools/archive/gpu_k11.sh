#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_lgr.py tests/test_procrustes.py -m gpu -q --timeout 300 > gpurun_out/test_lgr.log 2>&1
echo "test_lgr exit $?" | tee -a gpurun_out/summary.txt
tail -n 40 gpurun_out/test_lgr.log
timeout 300 python - > gpurun_out/lgr_timing.json 2> gpurun_out/lgr_timing.err <<'PY'
import json, sys
sys.argv = ['bench.py']
import bench, pkgpath
pkg = pkgpath.load_package()
print(json.dumps({s: bench.time_registration(pkg, s) for s in ('3dmatch', 'kitti')}, indent=1))
PY
echo "lgr_timing exit $?" | tee -a gpurun_out/summary.txt
cat gpurun_out/lgr_timing.json; tail -n 5 gpurun_out/lgr_timing.err
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/lgr_launches.csv python tools/lgr_once.py > gpurun_out/lgr_ncu.log 2>&1
python tools/ncu_launches.py gpurun_out/lgr_launches.csv 6
