#!/bin/bash
# tools/gpu_retry.sh [--gpus N] TIMEOUT_S 'command' -- runs `gpurun`, retrying while the pod answers busy (nothing is charged for those).
GPUS=""
if [ "$1" = "--gpus" ]; then GPUS="--gpus $2"; shift 2; fi
T=$1; shift
for attempt in $(seq 1 40); do
  /usr/local/graft/bin/gpurun $GPUS --timeout "$T" -- "$@" > /tmp/gpurun_last.txt 2>&1
  rc=$?
  if grep -q "status=transient\|nothing was charged" /tmp/gpurun_last.txt && [ $rc -ne 0 -o -z "$(grep -c 'rc=0' /tmp/gpurun_last.txt)" ] ; then
    if grep -q "status=transient" /tmp/gpurun_last.txt; then sleep 60; continue; fi
  fi
  if [ $rc -eq 3 ]; then sleep 60; continue; fi
  break
done
cat /tmp/gpurun_last.txt
exit $rc
