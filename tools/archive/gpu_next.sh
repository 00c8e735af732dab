#!/bin/bash
# GPU box: parity tests of the 'next' rows (K6, K7), their timings, and ncu evidence.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_next.py -m gpu -q --timeout 600 > gpurun_out/test_next.log 2>&1
echo "test_gpu_next exit $?" | tee -a gpurun_out/summary.txt
tail -n 25 gpurun_out/test_next.log
timeout 600 python tools/next_bench.py > gpurun_out/next_bench.json 2> gpurun_out/next_bench.err
echo "next_bench exit $?" | tee -a gpurun_out/summary.txt
cat gpurun_out/next_bench.json; tail -n 5 gpurun_out/next_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/next_launches.csv \
  python tools/next_bench.py --once > gpurun_out/ncu_next_launches.log 2>&1
echo "ncu launches exit $?" | tee -a gpurun_out/summary.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"log_optimal_transport|p2n_" -c 25 \
  -o gpurun_out/next_full python tools/next_bench.py --once > gpurun_out/ncu_next_full.log 2>&1
echo "ncu full exit $?" | tee -a gpurun_out/summary.txt
