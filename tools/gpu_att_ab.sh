# A/B of the attention kernel's knobs inside the pipelined step (tools/ablate_step.py --full-only)
run() { echo "== $W $*"; env "$@" timeout 120 python tools/ablate_step.py --full-only --workload $W --steps $S 2>&1 | tail -1; }
W=3dmatch_1024; S=60
run GEOTX_ATT=f16
W=3dmatch_4096; S=12
run GEOTX_ATT=f16
W=3dmatch_2048; S=30
run GEOTX_ATT=f16
W=kitti_2048; S=30
run GEOTX_ATT=f16
W=3dmatch_512; S=100
run GEOTX_ATT=f32
run GEOTX_ATT=f16
