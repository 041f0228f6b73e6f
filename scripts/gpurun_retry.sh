#!/bin/bash
# usage: scripts/gpurun_retry.sh <timeout-seconds> '<command>' [gpus]   -- retries while the pod answers busy/transient
T=$1; CMD=$2; G=${3:-1}
for i in $(seq 1 30); do
  if [ "$G" = "1" ]; then OUT=$(/usr/local/graft/bin/gpurun --timeout $T -- "$CMD" 2>&1); else OUT=$(/usr/local/graft/bin/gpurun --gpus $G --timeout $T -- "$CMD" 2>&1); fi
  if echo "$OUT" | grep -q "status=transient\|retry in a few minutes\|status=busy"; then sleep 90; continue; fi
  echo "$OUT"; exit 0
done
echo "$OUT"; exit 3
