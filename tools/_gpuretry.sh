#!/bin/bash
# usage: tools/_gpuretry.sh <gpus> <timeout-seconds> <command...>; retries while the pod is busy (exit 3)
G=$1; T=$2; shift; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --gpus $G --timeout $T -- "$@" > gpurun_out/.retry.log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then tail -25 gpurun_out/.retry.log; exit $rc; fi
  sleep 45
done
tail -5 gpurun_out/.retry.log; exit 3
