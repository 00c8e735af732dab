#!/bin/bash
# GPU box: sweep of the L2 slab size of the embedding -> first-block scores interleave (GEOTX_L2_SLAB_BYTES).
mkdir -p gpurun_out
for wl in ${SLAB_WORKLOADS:-3dmatch_1024 3dmatch_4096}; do
  for mb in ${SLAB_MB:-0 24 32 48 64 96}; do
    GEOTX_L2_SLAB_BYTES=$((mb << 20)) timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline \
      > gpurun_out/slab_${wl}_${mb}.json 2> gpurun_out/slab_${wl}_${mb}.err
    python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/slab_${wl}_${mb}.json').read().strip().splitlines()[-1])
    r = d['roofline']
    print('$wl slab ${mb} MB: %.2f pairs/s  %.3f ms/step  K2 %.0f TF/s over %d launches (share %.2f)' % (
        d['value'], d['ms_per_step'], r['achieved'], r['launches_timed'], r['share_of_step']))
except Exception as e:
    print('$wl slab ${mb}: failed', e)
PY
  done
done
