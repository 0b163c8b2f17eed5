#!/bin/bash
# gpurun with retries while the pod answers "busy" (status=transient: nothing is charged).
#   scripts/gpurun_retry.sh <timeout-seconds> '<command>'   [GPUS=N for multi-GPU calls]
t=$1; shift
for i in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun ${GPUS:+--gpus $GPUS} --timeout "$t" -- "$@" 2>&1)
  if echo "$out" | grep -q "status=transient"; then sleep 150; continue; fi
  echo "$out"; exit 0
done
echo "gave up: pod busy"; exit 3
