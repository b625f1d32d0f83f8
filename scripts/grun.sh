#!/bin/bash
# usage: scripts/grun.sh <log-name> <timeout-s> '<command>'  -- retries gpurun while the pod answers "transient"/busy
name=$1; to=$2; shift 2
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout $to -- "$@" > gpurun_out/$name.txt 2>&1
  if ! grep -q "status=transient\|exit code 3\|rc=3 " gpurun_out/$name.txt; then break; fi
  sleep 90
done
tail -3 gpurun_out/$name.txt
