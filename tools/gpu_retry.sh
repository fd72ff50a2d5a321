#!/bin/bash
# usage: tools/gpu_retry.sh <log> <timeout> <command...>   -- retries gpurun while the pod answers "busy"
log=$1; to=$2; shift 2
for i in 1 2 3 4 5 6 7 8 9 10; do
  /usr/local/graft/bin/gpurun --timeout $to -- "$*" > $log 2>&1
  grep -q "status=transient\|rc=3\|refused" $log || break
  sleep 90
done
