#!/bin/bash
# one --set full capture of the attention kernel K3h alone (n = 1024, no key split: the throughput schedule), a launch
# with scores and one without
mkdir -p gpurun_out
timeout 120 python tools/att_once.py 1024 0 > gpurun_out/att_plain.log 2>&1 && timeout 120 python tools/att_once.py 4096 0 >> gpurun_out/att_plain.log 2>&1 && timeout 120 python tools/att_once.py 1024 1 >> gpurun_out/att_plain.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on -k regex:attention_f16 -s 2 -c 2 -f -o gpurun_out/r02_k3h python tools/att_once.py 1024 0 > gpurun_out/att_ncu.log 2>&1
echo rc=$?; cat gpurun_out/att_plain.log; tail -2 gpurun_out/att_ncu.log
