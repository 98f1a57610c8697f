#!/bin/bash
# usage: tools/gpurun_retry.sh TIMEOUT 'command' [extra gpurun args]   -- retries while the pod answers transient/busy (rc 3)
T=$1; CMD=$2; shift 2; mkdir -p gpurun_out
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" --timeout $T -- "$CMD" > gpurun_out/.gpurun_retry.log 2>&1; rc=$?
  if ! grep -q "status=transient\|status=busy" gpurun_out/.gpurun_retry.log && [ $rc -ne 3 ]; then break; fi
  sleep 45
done
grep -v WARNING gpurun_out/.gpurun_retry.log | tail -${TAIL:-40}
