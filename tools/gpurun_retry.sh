#!/bin/bash
# gpurun with retries on exit code 3 (no box / slot free right now; nothing charged).
# usage: tools/gpurun_retry.sh <log> <gpurun args...>
log=$1; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  [[ $rc != 3 ]] && exit $rc
  sleep 120
done
exit 3
