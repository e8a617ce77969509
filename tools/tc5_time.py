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
orb.lib().orb_gpu_set_bruteforce_variant(variant, 2)
ts = []
for _ in range(6):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); parallel.match_bruteforce_sharded(q, db, 0); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = float(np.median(ts[1:]))
print("dbg=%s variant %d: %8.3f ms  %8.1f Gpairs/s" % (os.environ.get("ORB_GPU_TC5_DEBUG", "0"), variant, ms, 4096 * ndb / ms / 1e6), flush=True)
