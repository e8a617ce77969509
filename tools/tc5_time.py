#!/usr/bin/env python
"""Device time of the tcgen05 brute-force form (ORB_GPU_TC5_DEBUG selects what is skipped; timing only)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import c_orb_slam_b200 as orb
from c_orb_slam_b200 import parallel
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(11)
ndb = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
variant = int(sys.argv[2]) if len(sys.argv) > 2 else 3
db = torch.randint(0, 256, (ndb, 32), dtype=torch.uint8, device=dev, generator=g)
q = torch.randint(0, 256, (4096, 32), dtype=torch.uint8, device=dev, generator=g)
for v in (2, variant):
    orb.lib().orb_gpu_set_bruteforce_variant(v, 2)
    ts = []
    for _ in range(16):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); parallel.match_bruteforce_sharded(q, db, 0); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts = ts[1:]
    print("variant %d: min %8.3f ms median %8.3f ms max %8.3f  -> %8.1f Gpairs/s (min) %8.1f (median)" % (
        v, min(ts), float(np.median(ts)), max(ts), 4096 * ndb / min(ts) / 1e6, 4096 * ndb / float(np.median(ts)) / 1e6), flush=True)
