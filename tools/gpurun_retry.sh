#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout_s> <script> [gpus]   — retries while the pod answers "busy" (exit 3, nothing charged)
T=$1; S=$2; G=${3:-1}
for i in $(seq 1 12); do
  if [ "$G" = 1 ]; then /usr/local/graft/bin/gpurun --timeout "$T" -- "bash $S"; else /usr/local/graft/bin/gpurun --gpus "$G" --timeout "$T" -- "bash $S"; fi
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 150
done
exit 3
