#!/bin/bash
# usage: scripts/gpurun_retry.sh <logfile> <gpurun args...>   — retries while the pod answers "transient / busy" (exit 3: nothing charged)
log=$1; shift
for attempt in $(seq 1 30); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" "$log"; then exit $rc; fi
  sleep 45
done
exit 3
