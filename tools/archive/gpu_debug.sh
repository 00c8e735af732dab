#!/bin/bash
mkdir -p gpurun_out
for v in "one 256" "one 128" "pair 256" "pair 128"; do
  timeout 60 tools/bin/umma_selftest $v 2>&1 | tee -a gpurun_out/selftest.log
  echo "selftest $v exit $?" | tee -a gpurun_out/selftest.log
done
for a in "256 64 max" "128 64 max" "256 300 mean"; do
  CUDA_LAUNCH_BLOCKING=1 timeout 120 python tools/k2_debug.py $a 2>&1 | grep -v Warning | tail -12 | tee -a gpurun_out/k2_debug.log
  echo "k2_debug $a exit $?" | tee -a gpurun_out/k2_debug.log
done
timeout 300 python -m pytest tests/test_gpu_k1_knn.py -m gpu -q --timeout 200 2>&1 | tail -5 | tee gpurun_out/test_k1_knn.log
