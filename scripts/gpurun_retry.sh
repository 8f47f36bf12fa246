#!/bin/bash
# gpurun with retries while the pod answers "transient" (busy): scripts/gpurun_retry.sh <log> [gpurun options] -- '<command>'
LOG=$1; shift
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  grep -q "status=transient" "$LOG" || break
  sleep 45
done
tail -25 "$LOG" | cut -c1-220
