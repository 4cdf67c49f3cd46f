#!/bin/bash
# usage: dev/gpurun_retry.sh <log> <gpurun args...>   retries while the pod answers "busy" (exit 3)
log=$1; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1; rc=$?
  echo "exit $rc (attempt $i)" >> "$log"
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 100
done
exit 3
