#!/bin/bash
# gpurun with retries while the pod answers "busy / draining" (nothing is charged for those).
# usage: scripts/gpurun_retry.sh [gpurun options] -- 'command'
for attempt in $(seq 1 20); do
  out=$(/usr/local/graft/bin/gpurun "$@" 2>&1)
  echo "$out" | tail -25
  if echo "$out" | grep -q "status=transient\|retry in a few minutes\|no box or slot"; then
    sleep 90
    continue
  fi
  break
done
