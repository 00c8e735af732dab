#!/bin/bash
# one bench.py run per workload of the sweep (BASELINE.json configs[0..3]); lines land in gpurun_out/sweep_<tag>/
TAG=${TAG:-r02}
mkdir -p gpurun_out/sweep_$TAG
for wl in ${WORKLOADS:-3dmatch_512 3dmatch_1024 3dmatch_2048 3dmatch_4096 3dmatch_8192 3dmatch_16384 kitti_2048 modelnet_717}; do
  steps=20; case $wl in 3dmatch_4096) steps=10;; 3dmatch_8192) steps=5;; 3dmatch_16384) steps=3;; esac
  timeout 900 python bench.py --workload $wl --steps $steps --warmup 3 --no-sweep --no-cpu-baseline > gpurun_out/sweep_$TAG/$wl.json 2> gpurun_out/sweep_$TAG/$wl.err
  echo "$wl rc=$?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/sweep_$TAG/$wl.json') if l.startswith('{')][0])
    r=d['roofline'] or {}
    print('  value %.2f pairs/s  ms %.3f  e2e %.2f  K3r %.4f ms/launch frac_mma %.3f  mode %s  clocks %s'%(d['value'], d['ms_per_step'], d['e2e']['value'], r.get('avg_launch_ms',0), r.get('frac_of_mma_sync_peak',0), d['setup']['e_storage'][:8], d['clocks']))
    v=d.get('variants',{})
    for k in v: print('   ',k, v[k].get('value'), v[k].get('error'))
except Exception as e:
    print('  parse failed', e)
    import subprocess; print(open('gpurun_out/sweep_$TAG/$wl.err').read()[-600:])
PY
done
