#!/bin/bash
# Run one gpurun call, retrying while the pod answers "busy" (exit code 3: nothing charged).  usage: tools/gpu_retry.sh [gpurun flags] -- 'command'
for i in $(seq 1 40); do
    /usr/local/graft/bin/gpurun "$@"
    rc=$?
    if [ $rc -ne 3 ]; then exit $rc; fi
    sleep 75
done
exit 3
