#!/usr/bin/env python
"""Host-call latency of one TUM frame (configs[0] as written): orb_gpu_extract through ctypes, pinned staging and the
CUDA graph inside the library.  ORB_GPU_BLUR=simd for the byte-SIMD blur."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import c_orb_slam_b200 as orb
from c_orb_slam_b200 import synth
img = synth.make_frame(640, 480, 1000)
ex = orb.ORBextractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480)
for _ in range(20):
    ex(img)
for rep in range(3):
    t0 = time.perf_counter()
    for _ in range(300):
        ex.extract_batch(img[None])
    print("single frame host call: %.1f us" % ((time.perf_counter() - t0) / 300 * 1e6))
