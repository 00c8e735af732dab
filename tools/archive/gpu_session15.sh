#!/bin/bash
mkdir -p gpurun_out
TAG=${TAG:-v15}
timeout 300 python -m pytest tests/test_gpu_dense.py -m gpu -q -x 2>&1 | tail -15
timeout 500 python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/s15_tests.txt; tail -4 gpurun_out/s15_tests.txt
timeout 600 python bench.py --no-sweep --no-cpu-baseline > gpurun_out/r02_bench_$TAG.json 2> gpurun_out/r02_bench_$TAG.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02_bench_$TAG.err; head -c 200 gpurun_out/r02_bench_$TAG.json; echo
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r02_bench_$TAG.json') if l.startswith('{')][0])
for k,v in d['kernels'].items(): print(k, v['launches_per_step'], '%.1f us avg'%(v['avg_launch_ms']*1e3), '%.3f ms/step'%v['ms_per_step'])
print('serial', d['serial_ms_per_step'], 'e2e', d['e2e']['value'])
PY
