#!/bin/bash
# Retry wrapper around gpurun: exit code 3 means "no box or slot free right now, nothing charged" - wait and ask again.
#   tools/gpu_retry.sh <log file> [--gpus N] [--timeout S] -- '<command>'
LOG=$1; shift
for attempt in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 45
done
exit 3
