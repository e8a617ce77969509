#!/bin/bash
# host-chunk schedules of the host-pointer path (ORB_GPU_HOST_SCHEDULE = head / tail chunk sizes)
E2E_BW=1 python tools/e2e_sweep.py
for s in ${E2E_SCHEDULES:-"16,32,64,128/64,32,16" "16,32,64,96/48,32,16" "32,64,128/64,32" "16,48,96/64,32,16" "32,64/32" "16,32,64/32,16" "64/32,16" "32,96/64,32,16"}; do
  ORB_GPU_HOST_SCHEDULE=$s python tools/e2e_sweep.py | sed "s|^|sched $s: |"
done
ORB_GPU_TIMELINE=1 python tools/e2e_sweep.py 2>&1 | grep timeline | tail -16
ORB_GPU_HOST_SCHEDULE="16,32,64,128/64,32,16" ORB_GPU_TIMELINE=1 python tools/e2e_sweep.py 2>&1 | grep timeline | tail -8
