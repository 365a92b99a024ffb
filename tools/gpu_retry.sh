#!/bin/bash
# Run a command on the GPU pool, retrying while the pod answers busy / draining (exit 3 = nothing charged).
#   tools/gpu_retry.sh <timeout seconds> '<command>' [extra gpurun flags, e.g. --gpus 2]
T=$1; CMD=$2; shift 2
for i in $(seq 1 30); do
    /usr/local/graft/bin/gpurun --timeout "$T" "$@" -- "$CMD"
    rc=$?
    if [ $rc -ne 3 ] && ! grep -q '"status": *"transient"' gpurun_out/.last_call.json 2>/dev/null; then exit $rc; fi
    sleep 60
done
exit 3
