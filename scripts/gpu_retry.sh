#!/bin/bash
# usage: scripts/gpu_retry.sh LOGFILE TIMEOUT [--gpus N] -- 'command'   (retries while the pod answers busy / transient)
LOG=$1; shift; TO=$1; shift
for i in $(seq 1 30); do
    /usr/local/graft/bin/gpurun --timeout $TO "$@" > $LOG 2>&1
    rc=$?
    if grep -q "status=transient\|status=busy" $LOG || [ $rc -eq 3 ]; then sleep 45; continue; fi
    exit $rc
done
exit 3
