#!/bin/bash
# usage: gpu_retry.sh <logfile> <gpurun args...>   -- retries while the pod answers "transient"
log=$1; shift
for i in $(seq 1 12); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  if ! grep -q "status=transient" "$log"; then break; fi
  sleep 100
done
