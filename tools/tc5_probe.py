#!/usr/bin/env python
"""Parity and device time of the tcgen05 brute-force form (variant 3) against the mma.sync form (variant 2)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import c_orb_slam_b200 as orb
from c_orb_slam_b200 import parallel
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(11)


def run(variant, q, db):
    orb.lib().orb_gpu_set_bruteforce_variant(variant, 2)
    out = parallel.match_bruteforce_sharded(q, db, 0)
    torch.cuda.synchronize()
    return out


sizes = [(1, 1), (100, 300), (256, 256), (300, 1000), (1000, 5000), (2000, 70000), (4096, 1_000_000)]
if len(sys.argv) > 1:
    sizes.append((4096, int(sys.argv[1])))
for nq, ndb in sizes:
    db = torch.randint(0, 256, (ndb, 32), dtype=torch.uint8, device=dev, generator=g)
    q = torch.randint(0, 256, (nq, 32), dtype=torch.uint8, device=dev, generator=g)
    if ndb >= 300:  # plant near duplicates and exact ties
        db[ndb // 2] = q[0]; db[ndb // 3] = q[0]; db[7] = q[nq - 1]
    a = run(2, q, db).cpu().numpy()
    b = run(3, q, db).cpu().numpy()
    same = np.array_equal(a, b)
    print("nq=%d ndb=%d identical=%s" % (nq, ndb, same), flush=True)
    if not same:
        bad = np.argwhere(a != b)
        print(" first diffs:", bad[:8].tolist(), a[tuple(bad[0])], b[tuple(bad[0])])
        print(" a:", a[..., : 4] if a.ndim > 1 else a[:4]); print(" b:", b[..., :4] if b.ndim > 1 else b[:4])
        sys.exit(1)
    if ndb >= 1_000_000:
        for v in (2, 3):
            ts = []
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); run_out = None
                orb.lib().orb_gpu_set_bruteforce_variant(v, 2)
                parallel.match_bruteforce_sharded(q, db, 0); e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            ms = float(np.median(ts))
            print("  variant %d: %8.3f ms  %8.1f Gpairs/s" % (v, ms, nq * ndb / ms / 1e6), flush=True)
orb.lib().orb_gpu_set_bruteforce_variant(2, 2)
