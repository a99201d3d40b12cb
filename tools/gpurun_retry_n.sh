#!/bin/bash
# gpurun --gpus N with retries while no slot is free.  usage: tools/gpurun_retry_n.sh N LOG TIMEOUT 'command'
N=$1; LOG=$2; TO=$3; shift 3
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --gpus $N --timeout $TO -- "$@" > $LOG 2>&1
  RC=$?
  if [ $RC -ne 3 ] && ! grep -q "status=transient" $LOG; then exit $RC; fi
  sleep 60
done
exit 3
