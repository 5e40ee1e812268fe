#!/bin/bash
# usage: scripts/gpurun_retry.sh <logfile> <timeout_s> [--gpus N] <command>   retries while gpurun answers 3 (no slot free)
log=$1; shift; to=$1; shift
extra=""
if [ "$1" == "--gpus" ]; then extra="--gpus $2"; shift; shift; fi
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $to $extra -- "$@" > $log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
