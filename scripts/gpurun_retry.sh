#!/bin/bash
# gpurun with retries while the pod answers "busy" (nothing is charged for those): scripts/gpurun_retry.sh <timeout> '<command>'
T=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$T" -- "$@" > /tmp/gpurun_retry.log 2>&1
  if ! grep -q "status=transient" /tmp/gpurun_retry.log; then break; fi
  sleep 120
done
tail -150 /tmp/gpurun_retry.log
