#!/bin/bash
# gpu_retry.sh <timeout_s> <command...>: gpurun with retries while the pod answers "busy / draining" (nothing charged).
# GPUS=2|4|8 in the environment asks for that many GPUs of one box.
T=$1; shift
G=""; [ -n "$GPUS" ] && G="--gpus $GPUS"
for i in $(seq 1 40); do
  out=$(/usr/local/graft/bin/gpurun $G --timeout $T -- "$@" 2>&1)
  if echo "$out" | grep -q "status=transient\|status=busy\|nothing was charged"; then sleep 45; continue; fi
  echo "$out"; exit 0
done
echo "$out"; exit 3
