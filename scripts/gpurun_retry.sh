#!/bin/bash
# scripts/gpurun_retry.sh <log> <gpurun args...>: retry while the pool answers busy / transient (exit 3), up to ~40 min
log=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1; rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" "$log"; then exit $rc; fi
  sleep 60
done
exit 3
