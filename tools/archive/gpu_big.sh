#!/bin/bash
# GPU box: E-storage modes + the two large sweep points of BASELINE configs[3].
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_forward.py -m gpu -q --timeout 300 -k "regenerated" > gpurun_out/test_regen.log 2>&1
echo "test_regen exit $?" | tee -a gpurun_out/summary.txt
tail -n 8 gpurun_out/test_regen.log
for w in 3dmatch_8192 3dmatch_16384; do
  timeout 1200 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
  echo "bench $w exit $?" | tee -a gpurun_out/summary.txt
  tail -n 3 gpurun_out/bench_$w.err
  python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/bench_$w.json').read().strip().splitlines()[-1])
    print({k: d[k] for k in ('value', 'ms_per_step', 'pair_tflops', 'pair_tflops_min', 'pair_tensor_frac', 'clocks')}, d['config']['e_storage'], d['roofline']['achieved'], d['roofline']['launches_timed'], d['e2e']['value'])
except Exception as e:
    print('no line', e)
PY
done
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
