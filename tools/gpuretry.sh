#!/bin/bash
# Retry a gpurun call while the pod answers busy/transient.  usage: tools/gpuretry.sh <logfile> <timeout-seconds> '<command>'
LOG=$1; TO=$2; shift 2
for i in $(seq 1 12); do
  /usr/local/graft/bin/gpurun --timeout $TO -- "$@" > $LOG 2>&1
  if ! grep -q "status=transient\|status=busy" $LOG; then break; fi
  sleep 90
done
