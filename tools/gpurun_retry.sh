#!/bin/bash
# gpurun with retries while the pod answers "busy" (exit code 3 / transient); usage: tools/gpurun_retry.sh LOG [gpurun args...]
log=$1; shift
for attempt in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  if grep -q "status=transient" "$log" || [ $rc -eq 3 ]; then sleep 120; continue; fi
  break
done
echo "gpurun rc=$rc attempts=$attempt" >> "$log"
