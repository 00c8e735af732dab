#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_next.py -m gpu -q --timeout 600 > gpurun_out/test_next.log 2>&1
echo "test_gpu_next exit $?" | tee -a gpurun_out/summary.txt
tail -n 8 gpurun_out/test_next.log
timeout 600 python tools/next_bench.py > gpurun_out/next_bench.json 2> gpurun_out/next_bench.err
echo "next_bench exit $?" | tee -a gpurun_out/summary.txt
cat gpurun_out/next_bench.json; tail -n 5 gpurun_out/next_bench.err
