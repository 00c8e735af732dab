# A/B of attention-kernel knobs inside the pipelined step (tools/ablate_step.py --full-only) + their kernel tables
for gr in 1 2; do
  echo "=== GEOTX_ATT_GROUPS=$gr"
  for w in 3dmatch_1024 3dmatch_2048 kitti_2048 3dmatch_512; do
    S=60; [ $w = 3dmatch_2048 ] && S=30; [ $w = kitti_2048 ] && S=30
    echo -n "$w: "; GEOTX_ATT_GROUPS=$gr timeout 120 python tools/ablate_step.py --full-only --workload $w --steps $S 2>&1 | tail -1
  done
  GEOTX_ATT_GROUPS=$gr STEPS=20 bash tools/gpu_ktable.sh 3dmatch_1024 | grep -i "attention\|pairs"
done
