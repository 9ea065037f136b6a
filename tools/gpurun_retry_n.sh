#!/bin/bash
# usage: tools/gpurun_retry_n.sh <gpus> <timeout-seconds> '<command>'
N=$1; T=$2; shift; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --gpus "$N" --timeout "$T" -- "$@" > /tmp/gpurun_last_n.txt 2>&1
  rc=$?
  if grep -q "status=transient" /tmp/gpurun_last_n.txt || [ $rc -eq 3 ]; then sleep 120; continue; fi
  break
done
tail -40 /tmp/gpurun_last_n.txt
exit $rc
