#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout_s> <gpus> <command...>   — retries while the pod answers "busy" (exit 3), nothing is charged for those
t=$1; g=$2; shift 2
for attempt in $(seq 1 40); do
  if [ "$g" = "1" ]; then /usr/local/graft/bin/gpurun --timeout $t -- "$@"; else /usr/local/graft/bin/gpurun --gpus $g --timeout $t -- "$@"; fi
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  echo "[retry] attempt $attempt answered busy; sleeping 120 s"; sleep 120
done
exit 3
