#!/bin/bash
# gpu_retry.sh <timeout_s> [--gpus N] <command...> -- gpurun, retried while the pod answers "busy / draining"
T=$1; shift
G=""
if [ "$1" = "--gpus" ]; then G="--gpus $2"; shift 2; fi
for i in 1 2 3 4 5 6 7 8 9 10; do
  OUT=$(gpurun $G --timeout "$T" -- "$@" 2>&1)
  if echo "$OUT" | grep -q "status=transient\|nothing was charged"; then sleep 60; continue; fi
  echo "$OUT"; exit 0
done
echo "$OUT"; exit 3
