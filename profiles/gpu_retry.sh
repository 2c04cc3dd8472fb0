#!/bin/bash
# gpu_retry.sh <timeout_s> <command...> -- gpurun, retried while the pod answers "busy / draining"
T=$1; shift
for i in 1 2 3 4 5 6 7 8; do
  OUT=$(gpurun --timeout "$T" -- "$@" 2>&1)
  if echo "$OUT" | grep -q "status=transient\|nothing was charged"; then sleep 60; continue; fi
  echo "$OUT"; exit 0
done
echo "$OUT"; exit 3
