#!/usr/bin/env bash
# Runs gpurun and retries while the pod has no free slot (exit code 3):
#   tools/gpurun_retry.sh <log> <timeout-seconds> [--gpus N] '<command>'
log="$1"; shift
to="$1"; shift
extra=()
if [ "$1" = "--gpus" ]; then extra=(--gpus "$2"); shift 2; fi
for i in $(seq 1 30); do
  /usr/local/graft/bin/gpurun --timeout "$to" "${extra[@]}" -- "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ] && ! grep -q "status=transient" "$log"; then break; fi
  sleep 90
done
echo "finished rc=$rc" >> "$log"
