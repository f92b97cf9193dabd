#!/bin/bash
# developer helper: retry gpurun while the pod answers "no box / slot free right now" (exit code 3, nothing charged)
for i in $(seq 1 12); do
    /usr/local/graft/bin/gpurun "$@"
    rc=$?
    if [ $rc -ne 3 ]; then exit $rc; fi
    grep -q transient gpurun_out/.last_call.json 2>/dev/null
    sleep 90
done
exit 3
