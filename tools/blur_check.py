#!/usr/bin/env python
"""Blurred pyramid of the GPU extractor against cv2.GaussianBlur of its own (unblurred) levels, per frame and level,
with the location of any mismatch; then the blur stage time of a large batch (ORB_GPU_BLUR=simd selects the byte-SIMD
kernel for comparison).  usage: blur_check.py [shape ...]   (GPU box)"""
import os
import sys
import time

import cv2
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import c_orb_slam_b200 as orb
from c_orb_slam_b200 import synth

shapes = sys.argv[1:] or ["tum", "kitti", "euroc"]
bad_total = 0
for shape in shapes:
    W, H = synth.SHAPES[shape]
    for B in (1, 5):
        frames = synth.make_frames(W, H, B, seed0=300)
        ex = orb.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
        ex.extract_batch(frames)
        for f in range(B):
            for l in range(8):
                src = ex.pyramid_level(l, f)
                got = ex.pyramid_level(l, f, blurred=True)
                ref = cv2.GaussianBlur(src, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
                d = got != ref
                n = int(d.sum())
                bad_total += n
                if n:
                    ys, xs = np.nonzero(d)
                    print("%s B=%d frame %d level %d (%dx%d): %d mismatches, rows %d..%d cols %d..%d" %
                          (shape, B, f, l, src.shape[1], src.shape[0], n, ys.min(), ys.max(), xs.min(), xs.max()))
                    print("   rows with mismatches: %d   cols with mismatches: %d" % (len(set(ys)), len(set(xs))))
                    for y, x in list(zip(ys, xs))[:6]:
                        print("   (y=%d, x=%d) got %d ref %d" % (y, x, got[y, x], ref[y, x]))
        print("%s B=%d checked" % (shape, B), flush=True)
print("TOTAL mismatches:", bad_total)

# stage time of a resident batch
import torch
W, H = synth.SHAPES["tum"]
B = 256
frames = torch.from_numpy(synth.make_frames(W, H, B, seed0=1)).cuda()
ex = orb.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
cap = ex.capacity
kps = torch.empty((B, cap, 28), dtype=torch.uint8, device="cuda")
desc = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda")
n = torch.empty(B, dtype=torch.int32, device="cuda")
ex.set_stage_timing(True)
for it in range(4):
    ex.extract_batch_device(frames.data_ptr(), B, W, H, W, W * H, kps.data_ptr(), desc.data_ptr(), n.data_ptr())
    torch.cuda.synchronize()
    print("stage ms (B=%d):" % B, {k: round(v, 4) for k, v in ex.stage_ms().items()})
ex.set_stage_timing(False)
torch.cuda.synchronize()
for rep in range(2):
    t0 = time.perf_counter()
    for it in range(10):
        ex.extract_batch_device(frames.data_ptr(), B, W, H, W, W * H, kps.data_ptr(), desc.data_ptr(), n.data_ptr())
    torch.cuda.synchronize()
    print("whole pipeline, blur forked: %.3f ms per %d frames" % ((time.perf_counter() - t0) * 100, B))
