#!/bin/bash
# gpurun with retries while the pod has no free slot (exit code 3 = nothing charged).  usage: tools/gpurun_retry.sh LOG TIMEOUT 'command'
LOG=$1; TO=$2; shift 2
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout $TO -- "$@" > $LOG 2>&1
  RC=$?
  if [ $RC -ne 3 ] && ! grep -q "status=transient" $LOG; then exit $RC; fi
  sleep 45
done
exit 3
