#!/bin/bash
# tools/gpurun_retry.sh <timeout-seconds> '<command>' : gpurun that waits out "no box / pod draining" answers
t=$1; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout "$t" -- "$@" > /tmp/gpurun_last.log 2>&1
  rc=$?
  if grep -q "status=transient\|no box\|busy" /tmp/gpurun_last.log && [ $rc -ne 0 ] && ! grep -q "charged=[1-9]" /tmp/gpurun_last.log; then sleep 120; continue; fi
  break
done
tail -15 /tmp/gpurun_last.log
exit $rc
