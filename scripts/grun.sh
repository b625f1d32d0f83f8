#!/bin/bash
# usage: scripts/grun.sh <log-name> <timeout-s> '<command>'  -- retries gpurun while the pod answers "transient"/busy
# GPUS=N in the environment asks for N GPUs of one box
name=$1; to=$2; shift 2
extra=""; if [ -n "$GPUS" ]; then extra="--gpus $GPUS"; fi
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun $extra --timeout $to -- "$@" > gpurun_out/$name.txt 2>&1
  if ! grep -q "status=transient\|exit code 3\|rc=3 \|status=busy" gpurun_out/$name.txt; then break; fi
  sleep 90
done
tail -3 gpurun_out/$name.txt
