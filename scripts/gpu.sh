#!/bin/bash
# Retry wrapper around gpurun: exit code 3 (no box / slot free, nothing charged) is retried every 2 minutes.
#   scripts/gpu.sh [--gpus N] <timeout_s> '<command>' <logfile>
GPUS=""
if [ "$1" = "--gpus" ]; then GPUS="--gpus $2"; shift 2; fi
T=$1; CMD=$2; LOG=${3:-/tmp/gpurun.log}
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun $GPUS --timeout $T -- "$CMD" > $LOG 2>&1
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 120
done
exit 3
