#!/bin/bash
# GPU box: bench.py over every workload of the sweep (BASELINE.json configs + north_star sweep points).
mkdir -p gpurun_out
for wl in ${SWEEP:-3dmatch_512 3dmatch_1024 3dmatch_2048 3dmatch_4096 kitti_2048 modelnet_717}; do
  timeout 900 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/sweep_$wl.json 2> gpurun_out/sweep_$wl.err
  python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/sweep_$wl.json').read().strip().splitlines()[-1])
    r = d['roofline']
    v = (d.get('variants') or {}).get('other_embed_kernel') or {}
    print('$wl: %.1f pairs/s  %.3f ms  e2e %.1f  [%s] %s %.0f %s (%.2f of peak, share %.2f)  other kernel: %s  clocks %s %s' % (
        d['value'], d['ms_per_step'], d['e2e']['value'], d['config']['embed_kernel'].split()[0], r['kernel'].split()[0],
        r['achieved'], r['unit'], r['frac'], r['share_of_step'] or 0, v.get('value'), d['clocks']['sm_mhz'], d['clocks']['reasons']))
except Exception as e:
    print('$wl: failed', e)
PY
done
