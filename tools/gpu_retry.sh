#!/bin/bash
# usage: tools/gpu_retry.sh <timeout_s> '<command>'   -- retries gpurun while the pod answers "transient" (nothing charged)
T=$1; shift
for attempt in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun --timeout "$T" -- "$@" 2>&1)
  if echo "$out" | grep -q "status=transient"; then sleep 45; continue; fi
  echo "$out"; exit 0
done
echo "$out"; echo "gave up after 40 transient answers"; exit 3
