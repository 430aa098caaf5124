#!/bin/bash
# Retries a gpurun call while the pod answers "busy" (exit code 3: nothing charged).  usage: tools/gpurun_retry.sh LOG TIMEOUT 'command'
log=$1; limit=$2; shift 2
for attempt in $(seq 1 12); do
    /usr/local/graft/bin/gpurun --timeout "$limit" -- "$@" > "$log" 2>&1
    rc=$?
    [ $rc -ne 3 ] && exit $rc
    sleep 150
done
exit 3
