#!/bin/bash
# usage: scripts/gpurun_retry.sh <log> <gpurun args...>   - retries while the pool answers "busy" (exit 3 / transient)
log=$1; shift
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  if grep -q "status=transient" "$log" || [ $rc -eq 3 ]; then sleep 90; continue; fi
  exit $rc
done
exit 3
