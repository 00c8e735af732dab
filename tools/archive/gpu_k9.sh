#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_collate.py -m gpu -q --timeout 300 > gpurun_out/test_collate.log 2>&1
echo "test_collate exit $?" | tee -a gpurun_out/summary.txt
tail -n 12 gpurun_out/test_collate.log
timeout 600 python tools/next_bench.py > gpurun_out/next_bench.json 2> gpurun_out/next_bench.err
echo "next_bench exit $?" | tee -a gpurun_out/summary.txt
python -c "
import json; d=json.load(open('gpurun_out/next_bench.json'))
for k,v in d.items():
    if 'radius' in k or 'pyramid' in k: print(k, v)"; tail -n 5 gpurun_out/next_bench.err
