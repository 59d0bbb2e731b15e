#!/usr/bin/env bash
# gpurun with retries while the pod answers "busy" (exit 3): bash scripts/gpurun_retry.sh <log> <gpurun args...>
LOG=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 45
done
exit 3
