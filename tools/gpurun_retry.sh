#!/bin/bash
# Build-container helper: retry a gpurun call while the pod answers "busy" (exit code 3, nothing charged).
# usage: bash tools/gpurun_retry.sh TIMEOUT_S 'command' [gpurun args...]
T=$1; shift
CMD=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$T" "$@" -- "$CMD"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  echo "[retry $i] busy, sleeping 90 s"
  sleep 90
done
exit 3
