#!/usr/bin/env python
"""Device time of the brute-force matcher variants (4096 x 10M): 8-POPC, carry-save, mma.sync IMMA, tcgen05 kind::i8 (default)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import c_orb_slam_b200 as orb
from c_orb_slam_b200 import parallel
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(7)
ndb = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
db = torch.randint(0, 256, (ndb, 32), dtype=torch.uint8, device=dev, generator=g)
q = torch.randint(0, 256, (4096, 32), dtype=torch.uint8, device=dev, generator=g)
ref = None
for name, v in (("8-popc", (0, 2)), ("carry-save", (1, 2)), ("imma", (2, 2)), ("tcgen05", (3, 2))):
    orb.lib().orb_gpu_set_bruteforce_variant(*v)
    out = parallel.match_bruteforce_sharded(q, db, 0)
    ts = []
    for _ in range(7):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); out = parallel.match_bruteforce_sharded(q, db, 0); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    o = out.cpu().numpy()
    same = ref is None or np.array_equal(o, ref)
    ref = o if ref is None else ref
    ms = float(np.median(ts))
    print("%-10s %8.2f ms  %7.1f Gpairs/s  identical=%s" % (name, ms, 4096 * ndb / ms / 1e6, same))
