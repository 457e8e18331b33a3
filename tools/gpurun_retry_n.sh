#!/bin/bash
# usage: tools/gpurun_retry_n.sh <ngpus> <logfile> <timeout> <command...>   -- like gpurun_retry.sh with --gpus N
n=$1; shift; log=$1; shift; to=$1; shift
for i in $(seq 1 60); do
  /usr/local/graft/bin/gpurun --gpus $n --timeout $to -- "$@" > $log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc after $i tries" >> $log; exit $rc; fi
  sleep 200
done
exit 3
