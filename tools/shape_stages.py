#!/usr/bin/env python
"""Per-stage device times of one extraction batch for the three BASELINE.json frame shapes (device C ABI):
total per step (CUDA events around the call) next to the sum of the stage timers, which exclude the level-0
repack of frames whose rows are not 4-byte aligned (KITTI's 1241-px rows)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import c_orb_slam_b200 as orb
from c_orb_slam_b200 import synth
torch.cuda.set_stream(torch.cuda.Stream())   # a real stream: the C ABI launches on it, the events see the work
for shape, nfeat, B in (("tum", 1000, 512), ("kitti", 2000, 256), ("euroc", 1200, 512)):
    W, H = synth.SHAPES[shape]
    cap = nfeat + 24
    d_fr = torch.from_numpy(synth.make_frames_fast(W, H, B, seed0=2000)).cuda()
    ex = orb.ORBextractor(nfeat, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
    d_k = torch.empty((B, cap, 28), dtype=torch.uint8, device="cuda")
    d_d = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(B, dtype=torch.int32, device="cuda")
    d_s = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    step = lambda: ex.extract_batch_device(d_fr.data_ptr(), B, W, H, W, W * H, d_k.data_ptr(), d_d.data_ptr(),
                                           d_n.data_ptr(), d_s.data_ptr(), st)
    for _ in range(3): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): step()
    e1.record(); torch.cuda.synchronize()
    total = e0.elapsed_time(e1) / 10
    ex.set_stage_timing(True)
    acc = {}
    for _ in range(5):
        step()
        for k, v in ex.stage_ms().items(): acc[k] = acc.get(k, 0.0) + v / 5
    ex.set_stage_timing(False)
    px = sum(int(round(W / 1.2 ** l)) * int(round(H / 1.2 ** l)) for l in range(8))
    print("%-6s %dx%d x%d nfeat %d: %.3f ms/step = %.0f frames/s = %.1f Gpx/s; stages %s sum %.3f; kp/frame %.0f" % (
        shape, W, H, B, nfeat, total, B / total * 1e3, B * px / total / 1e6,
        {k: round(v, 3) for k, v in acc.items()}, sum(acc.values()), float(d_n.sum()) / B), flush=True)
