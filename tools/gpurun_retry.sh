#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout> '<command>'   — retries while the pod answers busy (exit 3)
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$1" -- "$2"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 45
done
exit 3
