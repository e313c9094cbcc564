#!/bin/bash
# gpurun with retries while the pod answers "no box or slot free right now" (exit code 3, nothing charged)
#   usage: scripts/gpurun_retry.sh <log file> <gpurun args...>
log=$1; shift
for attempt in $(seq 1 30); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 90
done
exit 3
