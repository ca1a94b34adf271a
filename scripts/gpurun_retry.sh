#!/bin/bash
# gpurun with retries while the pod answers "transient" (exit code 3: no box or slot free, nothing charged).
# usage: scripts/gpurun_retry.sh <log file> <gpurun args...>
LOG=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" "$LOG"; then exit $rc; fi
  sleep 120
done
exit 3
