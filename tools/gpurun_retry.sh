#!/bin/bash
# gpurun with retries while the pod answers "transient" (nothing charged): tools/gpurun_retry.sh <timeout_s> <log> <command...>
T=$1; LOG=$2; shift 2
for i in 1 2 3 4 5 6 7 8; do
  gpurun --timeout $T -- "$@" > $LOG 2>&1
  if grep -q "status=transient" $LOG; then sleep 120; continue; fi
  break
done
