#!/bin/bash
# usage: scripts/gpurun_retry.sh <timeout-seconds> <command...>   -- retries while the pod answers "busy" (nothing charged)
T=$1; shift
for k in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun --timeout "$T" "$@" 2>&1)
  if echo "$out" | grep -q "status=transient\|nothing was charged"; then sleep 150; continue; fi
  echo "$out"; exit 0
done
echo "gave up: pod busy"; exit 3
