#!/bin/bash
# A/B of the attention core inside the pipelined step (tools/ablate_step.py --full-only): K3h on half q | k | v (default)
# against K3b on fp32 (GEOTX_ATT=f32), and the cap of K3h's cluster split, per workload; then the kernel tables.
run() { echo -n "$W $*: "; env "$@" timeout 120 python tools/ablate_step.py --full-only --workload $W --steps $S 2>&1 | tail -1; }
for W in ${WORKLOADS:-3dmatch_512 3dmatch_1024 3dmatch_2048 3dmatch_4096 kitti_2048}; do
  S=60; case $W in 3dmatch_2048|kitti_2048) S=30;; 3dmatch_4096) S=12;; esac
  run GEOTX_ATT=f32
  run GEOTX_ATT=f16
done
STEPS=20 bash tools/gpu_ktable.sh 3dmatch_1024
