#!/usr/bin/env python
"""Times the host-pointer C-ABI call (orb_gpu_extract_batch) for one chunk size (ORB_GPU_HOST_CHUNK)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import c_orb_slam_b200 as orb
from c_orb_slam_b200 import synth
B, W, H = 512, 640, 480
cap = 1024
fr = torch.from_numpy(synth.make_frames_fast(W, H, B, seed0=1000)).pin_memory()
hk = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
hd = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
hn = torch.zeros(B, dtype=torch.int32).pin_memory()
ex = orb.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
L = orb.lib()
def step():
    rc = L.orb_gpu_extract_batch(ex._h, fr.data_ptr(), B, W, H, W, W * H, hk.data_ptr(), hd.data_ptr(), cap, hn.data_ptr())
    assert rc == 0
for _ in range(3): step()
t0 = time.perf_counter()
for _ in range(10): step()
dt = (time.perf_counter() - t0) / 10
print("chunk", os.environ.get("ORB_GPU_HOST_CHUNK", "default"), "ms", round(dt * 1e3, 3), "fps", round(B / dt), "kp", int(hn.sum()))
if os.environ.get("E2E_BW"):
    d = torch.empty_like(fr, device="cuda")
    for _ in range(2): d.copy_(fr, non_blocking=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5): d.copy_(fr, non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 5
    print("H2D pinned %.1f MB: %.3f ms = %.1f GB/s" % (fr.numel() / 1e6, dt * 1e3, fr.numel() / dt / 1e9))
    dk = torch.empty_like(hk, device="cuda")
    for _ in range(2): hk.copy_(dk, non_blocking=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5): hk.copy_(dk, non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 5
    print("D2H pinned %.1f MB: %.3f ms = %.1f GB/s" % (hk.numel() / 1e6, dt * 1e3, hk.numel() / dt / 1e9))
