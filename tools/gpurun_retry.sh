#!/bin/bash
# gpurun with retries while the pod answers "transient" (no box / slot free right now; nothing is charged)
# usage: tools/gpurun_retry.sh <logfile> <gpurun args...>
log=$1; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  if grep -q "status=transient" "$log"; then sleep 45; continue; fi
  break
done
