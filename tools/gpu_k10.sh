#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_collate.py -m gpu -q --timeout 300 -k grid > gpurun_out/test_collate.log 2>&1
echo "test_collate exit $?" | tee -a gpurun_out/summary.txt
tail -n 25 gpurun_out/test_collate.log
