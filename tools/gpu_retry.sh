#!/bin/bash
# usage: gpu_retry.sh <timeout> <cmd>  -- retry while the pod answers busy (exit 3 / transient)
for i in 1 2 3 4 5 6 7 8; do
  out=$(/usr/local/graft/bin/gpurun --timeout $1 -- "$2" 2>&1); rc=$?
  if echo "$out" | grep -q "status=transient"; then sleep 120; continue; fi
  echo "$out" | tail -${3:-25}; exit $rc
done
echo "gave up: pod busy"
