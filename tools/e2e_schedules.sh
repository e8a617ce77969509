#!/bin/bash
# host-chunk schedules of the host-pointer path (ORB_GPU_HOST_SCHEDULE = head / tail chunk sizes)
E2E_BW=1 python tools/e2e_sweep.py
for s in "8,8,16,32" "8,16,32/16,8" "16,32/16" "8,24,32/16,8,8" "32/16,8,8" "16" "24" "48" "16,48/16"; do
  ORB_GPU_HOST_SCHEDULE=$s python tools/e2e_sweep.py | sed "s|^|sched $s: |"
done
