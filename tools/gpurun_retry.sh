#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout-seconds> [--gpus N] -- '<command>' : retries while the pod answers "busy/transient" (nothing charged)
T=$1; shift
for i in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun --timeout $T "$@" 2>&1); rc=$?
  echo "$out" | tail -25
  if echo "$out" | grep -q "status=transient\|nothing was charged"; then sleep 120; continue; fi
  exit $rc
done
