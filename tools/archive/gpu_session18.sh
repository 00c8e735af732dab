#!/bin/bash
mkdir -p gpurun_out
for f in 2 3 4 5 6; do
GEOTX_INFLIGHT=$f GEOTX_PIPE_DEPTH=$f timeout 300 python bench.py --no-sweep --no-cpu-baseline > gpurun_out/inflight_$f.json 2> gpurun_out/inflight_$f.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/inflight_$f.json') if l.startswith('{')][0])
print('inflight $f: value %.1f  e2e %.1f  ms %.4f'%(d['value'], d['e2e']['value'], d['ms_per_step']))
PY
done
