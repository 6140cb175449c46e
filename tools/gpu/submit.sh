#!/bin/bash
# submit.sh <timeout-seconds> <log> <command...>: gpurun with retries while the pod answers "busy" (exit 3, nothing charged)
T=$1; LOG=$2; shift 2
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout $T "$@" > $LOG 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" $LOG; then echo "gpurun rc=$rc after $i tries" >> $LOG; exit $rc; fi
  sleep 120
done
echo "gave up" >> $LOG; exit 3
