#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_tail_chain.py -m gpu -q --timeout 300 > gpurun_out/test_tail_chain.log 2>&1
echo "test_tail_chain exit $?" | tee -a gpurun_out/summary.txt
tail -n 30 gpurun_out/test_tail_chain.log
timeout 300 python tools/tail_bench.py > gpurun_out/tail_bench.json 2> gpurun_out/tail_bench.err
echo "tail_bench exit $?"; cat gpurun_out/tail_bench.json; tail -n 5 gpurun_out/tail_bench.err
