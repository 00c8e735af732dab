#!/bin/bash
# Runs on the GPU box (via gpurun): every -m gpu test file in its own process (a trap in one kernel must not poison
# the CUDA context of the others), then smoke() and a short bench.  Logs land in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
for f in gpu_k1_knn gpu_k3_attention gpu_k2_embed gpu_forward gpu_k5_matching gpu_next procrustes collate lgr tail_chain patch_scores; do
  timeout 600 python -m pytest tests/test_$f.py -m gpu -q -s --timeout 300 > gpurun_out/test_$f.log 2>&1
  echo "test_$f exit $?" | tee -a gpurun_out/summary.txt
  tail -n 15 gpurun_out/test_$f.log
done
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "smoke exit $?" | tee -a gpurun_out/summary.txt
tail -n 5 gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2>&1
echo "bench exit $?" | tee -a gpurun_out/summary.txt
tail -n 3 gpurun_out/bench.log
# launch list of the same bench command (ncu serialises and runs cold: shares, not absolute times)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
echo "ncu launches exit $?" | tee -a gpurun_out/summary.txt
