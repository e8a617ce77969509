import os, sys
sys.path.insert(0, "/root/repo")
import numpy as np, torch
import c_orb_slam_b200 as orb
from c_orb_slam_b200 import synth
W, H = synth.SHAPES["tum"]; B = 256
frames = torch.from_numpy(synth.make_frames(W, H, B, seed0=1)).cuda()
for min_th in (7, 20):
    ex = orb.ORBextractor(1000, 1.2, 8, 20, min_th, max_width=W, max_height=H, max_batch=B)
    cap = ex.capacity
    kps = torch.empty((B, cap, 28), dtype=torch.uint8, device="cuda"); desc = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda"); n = torch.empty(B, dtype=torch.int32, device="cuda")
    ex.set_stage_timing(True)
    for it in range(4):
        ex.extract_batch_device(frames.data_ptr(), B, W, H, W, W * H, kps.data_ptr(), desc.data_ptr(), n.data_ptr())
        torch.cuda.synchronize()
    print("minTh", min_th, {k: round(v, 4) for k, v in ex.stage_ms().items()}, "kp/frame", float(n.float().mean()))
