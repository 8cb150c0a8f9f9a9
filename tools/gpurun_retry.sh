#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout_s> '<command>'  -- retries while the pod answers "transient" (busy, nothing charged)
T=$1; shift
for i in $(seq 1 200); do
  out=$(/usr/local/graft/bin/gpurun --timeout "$T" -- "$@" 2>&1)
  if echo "$out" | grep -q "status=transient"; then sleep 3; continue; fi
  echo "$out" | tail -120
  exit 0
done
echo "gave up: pod busy"
