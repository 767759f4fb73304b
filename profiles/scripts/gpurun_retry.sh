#!/bin/bash
# gpurun with retries while the pod answers "transient" (exit 3: no box or slot free, nothing charged).
# usage: gpurun_retry.sh <timeout-seconds> '<command>'   (extra leading args such as "--gpus 2" go in GPURUN_ARGS)
t=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun $GPURUN_ARGS --timeout "$t" -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 60
done
exit 3
