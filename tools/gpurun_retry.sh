#!/bin/bash
# tools/gpurun_retry.sh [--gpus N] TIMEOUT 'command' -- retries gpurun every 2 minutes while the pod answers "transient" (no slot, nothing charged)
GP=""
if [ "$1" = "--gpus" ]; then GP="--gpus $2"; shift 2; fi
T=$1; shift
for i in $(seq 1 60); do
  out=$(/usr/local/graft/bin/gpurun $GP --timeout "$T" -- "$@" 2>&1)
  if echo "$out" | grep -q "status=transient"; then sleep 120; continue; fi
  echo "$out"; exit 0
done
echo "gpurun_retry: gave up after 60 attempts"; exit 3
