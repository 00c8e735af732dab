#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_procrustes.py -m gpu -q --timeout 300 > gpurun_out/test_procrustes.log 2>&1
echo "test_procrustes exit $?" | tee -a gpurun_out/summary.txt
tail -n 12 gpurun_out/test_procrustes.log
timeout 600 python tools/next_bench.py > gpurun_out/next_bench.json 2> gpurun_out/next_bench.err
echo "next_bench exit $?" | tee -a gpurun_out/summary.txt
cat gpurun_out/next_bench.json; tail -n 5 gpurun_out/next_bench.err
