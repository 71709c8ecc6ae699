#!/bin/bash
# usage: scripts/gpu_retry.sh <log> <timeout> [--gpus N] -- cmd     (retries gpurun while the pod answers busy / transient)
log=$1; shift; to=$1; shift
for i in $(seq 1 60); do
  /usr/local/graft/bin/gpurun --timeout $to "$@" > $log 2>&1; rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" $log; then echo "attempt $i rc=$rc"; break; fi
  sleep 120
done
tail -80 $log
