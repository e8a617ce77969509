#!/bin/bash
# usage: tools/gpu_retry.sh <log> <gpurun args...> : retries a gpurun call while the pod answers busy (exit code 3 / transient)
LOG=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  if ! grep -q "status=transient" "$LOG"; then break; fi
  sleep 90
done
echo "gpu_retry done after $i tries" >> "$LOG"
