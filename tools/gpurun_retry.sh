#!/bin/bash
# usage: tools/gpurun_retry.sh <log> <timeout-seconds> <command...>   -- retries while the pool answers busy (exit 3)
log=$1; shift; to=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$to" -- "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -qi "transient\|draining" "$log"; then echo "gpurun rc=$rc (attempt $i)"; exit $rc; fi
  sleep 45
done
echo "gave up"; exit 3
