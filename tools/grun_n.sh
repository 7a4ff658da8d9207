#!/bin/bash
# tools/grun_n.sh <gpus> <timeout-seconds> <log-name> <command...>: gpurun --gpus N with retries while busy
N=$1; T=$2; NAME=$3; shift 3
mkdir -p gpurun_out
for attempt in 1 2 3 4 5 6 7 8; do
  /usr/local/graft/bin/gpurun --gpus "$N" --timeout "$T" -- "$@" > "gpurun_out/${NAME}_call.log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
