#!/bin/bash
# gpu_retry.sh <log> <timeout-s> <command...>: gpurun with retries while the pod answers "busy" (exit 3)
LOG=$1; TMO=$2; shift 2
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $TMO -- "$@" > $LOG 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" $LOG; then echo "gpurun rc=$rc" >> $LOG; exit $rc; fi
  sleep 60
done
echo "gave up" >> $LOG
