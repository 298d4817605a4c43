#!/bin/bash
# usage: scripts/gpurun_retry.sh LOGFILE [gpurun args...] : retries while the pod answers busy / transient (nothing is charged then)
log=$1; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  if grep -q "status=transient\|answers busy\|status=busy" "$log" || [ $rc -eq 3 ]; then sleep 90; continue; fi
  break
done
echo "retry-wrapper done rc=$rc attempts=$i" >> "$log"
