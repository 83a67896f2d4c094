#!/usr/bin/env bash
# Runs gpurun and retries while the pod has no free slot (exit code 3):
#   tools/gpurun_retry.sh <log> <timeout-seconds> '<command>'
log="$1"; shift
to="$1"; shift
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout "$to" -- "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" "$log"; then break; fi
  sleep 90
done
echo "finished rc=$rc" >> "$log"
