#!/bin/bash
# usage: tools/gpu_retry.sh <tag> <timeout_s> '<command>'  -- runs gpurun, retrying while the pod answers "transient"/busy
tag=$1; to=$2; shift 2
mkdir -p gpurun_out
for i in $(seq 1 12); do
  /usr/local/graft/bin/gpurun --timeout "$to" -- "$@" > gpurun_out/${tag}_call.log 2>&1
  rc=$?
  if grep -q "status=transient\|status=busy" gpurun_out/${tag}_call.log || [ $rc -eq 3 ]; then sleep 60; continue; fi
  break
done
echo "done rc=$rc" >> gpurun_out/${tag}_call.log
