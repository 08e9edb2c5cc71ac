#!/bin/bash
# gpurun with retries while the pod answers "busy" (exit 3 / status=transient: nothing charged).
#   tools/gpurun_retry.sh <log> <timeout-seconds> [--gpus N] -- '<command>'
LOG=$1; TMO=$2; shift 2
for attempt in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $TMO "$@" > $LOG 2>&1
  if grep -q "status=transient" $LOG || grep -q "status=busy" $LOG; then sleep 90; continue; fi
  break
done
tail -40 $LOG
