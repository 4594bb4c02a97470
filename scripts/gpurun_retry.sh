#!/bin/bash
# usage: scripts/gpurun_retry.sh <logfile> <timeout> <gpus> '<command>' : retries while the pool answers "transient"
log=$1; to=$2; gpus=$3; cmd=$4
for i in $(seq 1 30); do
  if [ "$gpus" = "1" ]; then /usr/local/graft/bin/gpurun --timeout $to -- "$cmd" > $log 2>&1; rc=$?
  else /usr/local/graft/bin/gpurun --gpus $gpus --timeout $to -- "$cmd" > $log 2>&1; rc=$?; fi
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 45
done
exit 3
