#!/usr/bin/env python
"""Device time of the default brute-force matcher (4096 queries) over database sizes: shows the fixed cost per
call that limits strong scaling when the database is sharded over 2/4/8 GPUs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import c_orb_slam_b200 as orb
from c_orb_slam_b200 import parallel
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(11)
db_all = torch.randint(0, 256, (10_000_000, 32), dtype=torch.uint8, device=dev, generator=g)
q = torch.randint(0, 256, (4096, 32), dtype=torch.uint8, device=dev, generator=g)
sizes = [int(a) for a in sys.argv[1:]] or [10_000_000, 5_000_000, 2_500_000, 1_250_000, 625_000, 100_000]
print('ORB_GPU_BF_WAVES =', os.environ.get('ORB_GPU_BF_WAVES'))
for ndb in sizes:
    db = db_all[:ndb]
    ts = []
    for _ in range(12):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); parallel.match_bruteforce_sharded(q, db, 0); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts = ts[2:]
    print("ndb %9d: min %7.3f ms median %7.3f ms -> %7.1f Gpairs/s; ideal from 10M rate would be %7.3f ms" % (
        ndb, min(ts), float(np.median(ts)), 4096 * ndb / float(np.median(ts)) / 1e6, 5.556 * ndb / 1e7), flush=True)
