#!/bin/bash
# GPU box: ncu evidence for the 'next' rows (K6, K7, K8): launch list + one full capture per kernel.
mkdir -p gpurun_out
timeout 600 python tools/next_bench.py --once > gpurun_out/next_once.log 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/next_launches.csv \
  python tools/next_bench.py --once > gpurun_out/ncu_next_launches.log 2>&1
echo "ncu launches exit $?" | tee -a gpurun_out/summary.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"log_optimal_transport|p2n_|weighted_procrustes" -c 40 \
  -f -o gpurun_out/next_full python tools/next_bench.py --once > gpurun_out/ncu_next_full.log 2>&1
echo "ncu full exit $?" | tee -a gpurun_out/summary.txt
