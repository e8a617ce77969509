#!/usr/bin/env python
"""Where a single TUM frame's host call goes: the device time of each stage at batch 1 (CUDA events between the
launches, graph off) beside the whole host call (graph on).  usage: latency_stages.py   (GPU box)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import c_orb_slam_b200 as orb
from c_orb_slam_b200 import synth

W, H = synth.SHAPES["tum"]
img = synth.make_frame(W, H, 1000)
ex = orb.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H)
for _ in range(20):
    ex(img)
ts = []
for rep in range(5):
    t0 = time.perf_counter()
    for _ in range(300):
        ex.extract_batch(img[None])
    ts.append((time.perf_counter() - t0) / 300 * 1e6)
print("host call, one frame: %.1f us (best of 5 x 300)" % min(ts))
frames = torch.from_numpy(img[None].copy()).cuda()
cap = ex.capacity
kps = torch.empty((1, cap, 28), dtype=torch.uint8, device="cuda")
desc = torch.empty((1, cap, 32), dtype=torch.uint8, device="cuda")
n = torch.empty(1, dtype=torch.int32, device="cuda")
ex.set_stage_timing(True)
acc = {}
for it in range(20):
    ex.extract_batch_device(frames.data_ptr(), 1, W, H, W, W * H, kps.data_ptr(), desc.data_ptr(), n.data_ptr())
    torch.cuda.synchronize()
    if it >= 5:
        for k, v in ex.stage_ms().items():
            acc.setdefault(k, []).append(v * 1e3)
ex.set_stage_timing(False)
print("device time per stage at batch 1 (us, median):", {k: round(float(np.median(v)), 1) for k, v in acc.items()},
      "sum %.1f" % sum(float(np.median(v)) for v in acc.values()))
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for it in range(200):
    ex.extract_batch_device(frames.data_ptr(), 1, W, H, W, W * H, kps.data_ptr(), desc.data_ptr(), n.data_ptr())
e1.record()
torch.cuda.synchronize()
print("resident frame, back-to-back device calls: %.1f us per frame" % (e0.elapsed_time(e1) / 200 * 1e3))
