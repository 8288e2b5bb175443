#!/bin/bash
# usage: tools/gpu_retry_n.sh <tag> <ngpus> <timeout_s> '<command>'
tag=$1; n=$2; to=$3; shift 3
mkdir -p gpurun_out
for i in $(seq 1 12); do
  /usr/local/graft/bin/gpurun --gpus "$n" --timeout "$to" -- "$@" > gpurun_out/${tag}_call.log 2>&1
  rc=$?
  if grep -q "status=transient\|status=busy" gpurun_out/${tag}_call.log || [ $rc -eq 3 ]; then sleep 90; continue; fi
  break
done
echo "done rc=$rc" >> gpurun_out/${tag}_call.log
