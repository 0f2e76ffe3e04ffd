#!/bin/bash
# gpurun with retries while the pod answers "busy / transient" (exit code 3): scripts/gpurun_retry.sh LOG TIMEOUT [--gpus N] -- 'command'
LOG=$1; TMO=$2; shift 2
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout $TMO "$@" > $LOG 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" $LOG; then break; fi
  sleep 60
done
echo "gpurun_retry done rc=$rc try=$i" >> $LOG
