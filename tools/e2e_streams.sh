#!/bin/bash
# compute streams x chunk size of the host-pointer path
for st in 4 6 8; do for ch in 16 24 32 48; do
  ORB_GPU_HOST_STREAMS=$st ORB_GPU_HOST_CHUNK=$ch python tools/e2e_sweep.py | sed "s|^|streams $st: |"
done; done
ORB_GPU_TIMELINE=1 python tools/e2e_sweep.py 2>&1 | grep timeline | tail -16
python -m pytest tests/test_gpu_extractor.py -m gpu -x -q 2>&1 | tail -2
