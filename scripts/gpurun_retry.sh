#!/usr/bin/env bash
# Retry a gpurun call while the pod answers "busy" (exit 3). Usage: scripts/gpurun_retry.sh TIMEOUT [--gpus N] -- 'command'
T="$1"; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$T" "$@" > gpurun_out/.retry_last.log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then cat gpurun_out/.retry_last.log | tail -60; exit $rc; fi
  sleep 150
done
echo "gave up after 40 busy answers"; exit 3
