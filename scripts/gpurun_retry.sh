#!/bin/bash
# gpurun with retries while the pod has no free GPU slot (exit code 3 = nothing charged): scripts/gpurun_retry.sh <gpurun args...>
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@"
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 90
done
exit 3
