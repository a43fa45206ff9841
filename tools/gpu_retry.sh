#!/bin/bash
# gpu_retry.sh <timeout> <log> <command...>: call gpurun, retrying while the pod answers busy (rc 3)
T=$1; LOG=$2; shift 2
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $T -- "$@" > $LOG 2>&1; rc=$?
  if grep -q "status=transient" $LOG; then sleep 45; continue; fi
  exit $rc
done
exit 3
