#!/bin/bash
# usage: retry_gpurun.sh LOG TIMEOUT 'command'   -- retries while the pod answers busy/transient
log=$1; to=$2; shift 2
for i in 1 2 3 4 5 6 7 8; do
  gpurun --timeout $to -- "$@" > $log 2>&1
  if grep -q "status=transient\|exit code 3\|rc=3" $log || [ $? -eq 3 ]; then sleep 60; continue; fi
  break
done
