#!/bin/bash
# gpurun with retries while the pod answers "busy" (exit code 3 / status=transient): tools/gpurun_retry.sh <timeout_s> '<command>'
T=$1; shift
for i in $(seq 1 30); do
  out=$(/usr/local/graft/bin/gpurun --timeout "$T" -- "$@" 2>&1); rc=$?
  if echo "$out" | grep -q "status=transient"; then sleep 150; continue; fi
  echo "$out"; exit $rc
done
echo "gpurun_retry: gave up"; exit 3
