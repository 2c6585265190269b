#!/bin/bash
# Development helper: run a gpurun call, retrying while the pod answers "transient" (nothing charged).
#   tools/gpurun_retry.sh TIMEOUT 'command'
for i in 1 2 3 4 5 6 7 8; do
    out=$(/usr/local/graft/bin/gpurun --timeout "$1" -- "$2" 2>&1)
    echo "$out" | tail -25
    echo "$out" | grep -q "status=transient" || exit 0
    sleep 120
done
