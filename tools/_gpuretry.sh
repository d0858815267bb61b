#!/bin/bash
# usage: tools/_gpuretry.sh <timeout-seconds> <command...>; retries while the pod is busy (exit 3)
T=$1; shift
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun --timeout $T -- "$@" > gpurun_out/.retry.log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then tail -15 gpurun_out/.retry.log; exit $rc; fi
  sleep 45
done
tail -5 gpurun_out/.retry.log; exit 3
