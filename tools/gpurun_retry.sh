#!/bin/bash
# usage: [GPUS=2] tools/gpurun_retry.sh <timeout-seconds> '<command>'   (retries while the pod answers busy / transient)
T=$1; shift
for attempt in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun ${GPUS:+--gpus $GPUS} --timeout "$T" -- "$@" 2>&1)
  if echo "$out" | grep -q "status=transient\|rc=3\|no box\|retry in a few minutes"; then
    echo "[retry $attempt] busy; sleeping 60 s" >&2
    sleep 60
    continue
  fi
  echo "$out"
  exit 0
done
echo "$out"
exit 3
