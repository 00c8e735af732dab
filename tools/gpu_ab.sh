#!/bin/bash
# A/B of variant libraries on the bench line: VARIANTS="name ..." (tools/bin/<name>.so), shipped library first
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 400 python bench.py --no-sweep --no-cpu-baseline --steps ${STEPS:-40} --workload ${WL:-3dmatch_1024} > gpurun_out/ab_$tag.json 2> gpurun_out/ab_$tag.err
python - $tag <<'PY'
import json,sys
d=json.loads([l for l in open(f'gpurun_out/ab_{sys.argv[1]}.json') if l.startswith('{')][0])
k=d['kernels']
print('%-10s value %.1f e2e %.1f ms %.4f serial %.3f | '%(sys.argv[1], d['value'], d['e2e']['value'], d['ms_per_step'], d['serial_ms_per_step'])+' '.join('%s %.1f'%(n.split()[0], v['avg_launch_ms']*1e3) for n,v in k.items()))
PY
}
run shipped A=1
for v in $VARIANTS; do run $v GEOTX_LIB=$PWD/tools/bin/$v.so; done
run shipped2 A=1
