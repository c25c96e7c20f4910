#!/bin/bash
# gpurun with retries on "no slot right now" (exit 3): tools/gpurun_retry.sh LOG [gpurun args...]
log=$1; shift
for i in $(seq 1 12); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  [ $rc -ne 3 ] && exit $rc
  sleep 120
done
exit 3
