#!/bin/bash
# usage: scratch/gpurun_retry.sh <timeout> '<command>'  -- retries while the pod answers busy / transient (nothing charged)
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun --timeout "$1" -- "$2" > /tmp/gpurun_try.log 2>&1
  rc=$?
  if grep -q "status=transient\|no box\|busy" /tmp/gpurun_try.log && ! grep -q "status=ok" /tmp/gpurun_try.log; then sleep 120; continue; fi
  break
done
tail -n 12 /tmp/gpurun_try.log
exit $rc
