#!/bin/bash
# usage: tools/gpurun_retry.sh <logfile> <gpurun args...>   -- retries while gpurun answers "no box free" (exit 3)
log=$1; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1; rc=$?
  echo "exit $rc (attempt $i)" >> "$log"
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 150
done
