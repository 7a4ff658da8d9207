#!/bin/bash
# tools/grun.sh <timeout-seconds> <log-name> <command...>: gpurun with retries while the pod answers "busy" (exit 3)
T=$1; NAME=$2; shift 2
mkdir -p gpurun_out
for attempt in 1 2 3 4 5 6 7 8 9 10; do
  /usr/local/graft/bin/gpurun --timeout "$T" -- "$@" > "gpurun_out/${NAME}_call.log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
