#!/bin/bash
# usage: scripts/gpurun_retry.sh <timeout-seconds> [--gpus N] '<command>'   (retries while the pod has no free GPU slot)
t=$1; shift
extra=()
if [ "$1" = "--gpus" ]; then extra=(--gpus "$2"); shift 2; fi
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$t" "${extra[@]}" -- "$@" > /tmp/gpurun_last.log 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" /tmp/gpurun_last.log; then break; fi
  sleep 120
done
tail -60 /tmp/gpurun_last.log
exit $rc
