#!/bin/bash
# B200_PROFILING.md launch list of one forward of the bench workload: every launch with its device time
mkdir -p gpurun_out
timeout 200 python bench.py --launch-list --steps 2 > gpurun_out/ll_plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02_launches_3dmatch_1024.csv python bench.py --launch-list --steps 2 > gpurun_out/ll_ncu.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/r02_launches_3dmatch_1024.csv
