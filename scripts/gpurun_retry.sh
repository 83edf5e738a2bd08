#!/bin/bash
# gpurun with retries while the pod answers "busy / draining" (exit code 3 or status=transient): nothing is charged for those.
# usage: scripts/gpurun_retry.sh LOGFILE [gpurun args...]
LOG=$1; shift
for i in $(seq 1 12); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  rc=$?
  if ! grep -q "status=transient" "$LOG"; then exit $rc; fi
  sleep 150
done
exit 3
