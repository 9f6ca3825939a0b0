#!/bin/bash
# usage: [GPUS=8] tools/gpurun_retry.sh <logfile> <timeout_s> <command string>
# Retries while gpurun answers 3 (no box / pod busy; nothing charged).  The tree is snapshotted when a box is
# acquired, so keep /root/repo consistent while this waits.
log=$1; to=$2; shift 2
for i in $(seq 1 60); do
  /usr/local/graft/bin/gpurun ${GPUS:+--gpus $GPUS} --timeout $to -- "$@" > $log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "[retry] done rc=$rc after $i tries" >> $log; exit $rc; fi
  sleep 90
done
echo "[retry] gave up" >> $log; exit 3
