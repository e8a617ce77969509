#!/usr/bin/env python
"""Under torchrun (one process per GPU): the peer-memory exchange kernel against the NCCL all_gather + merge form of
the sharded brute-force matcher -- identical results, and the device time of both.
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tools/exchange_check.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import c_orb_slam_b200 as orb
from c_orb_slam_b200 import parallel
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
torch.cuda.set_stream(torch.cuda.Stream(device=dev))
dist.init_process_group("nccl", device_id=dev)
nq = 4096
ex = parallel.Top2Exchange(nq, device=local)
ok = True
for ndb in (1000, 200_000, 10_000_000):
    lo, hi = parallel.shard_range(ndb, rank, world)
    g = torch.Generator(device=dev); g.manual_seed(7 + rank)
    db = torch.randint(0, 256, (hi - lo, 32), dtype=torch.uint8, device=dev, generator=g)
    if hi > lo:
        db[:: max(1, (hi - lo) // 50)] = 0x5a          # equal rows in every shard: ties across ranks
    g.manual_seed(99)
    q = torch.randint(0, 256, (nq, 32), dtype=torch.uint8, device=dev, generator=g)
    q[:64] = 0x5a
    a = parallel.match_bruteforce_sharded(q, db, lo)                   # NCCL form
    b = parallel.match_bruteforce_sharded(q, db, lo, exchange=ex)      # peer-memory kernel
    same = bool(torch.equal(a, b))
    t = torch.tensor([int(same)], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MIN)
    ok &= bool(t.item())
    res = {}
    for name, e in (("nccl", None), ("peer", ex)):
        ts = []
        for _ in range(12):
            dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); parallel.match_bruteforce_sharded(q, db, lo, exchange=e); e1.record(); torch.cuda.synchronize()
            v = torch.tensor([e0.elapsed_time(e1)], device=dev); dist.all_reduce(v, op=dist.ReduceOp.MAX)
            ts.append(float(v.item()))
        res[name] = float(np.median(ts[2:]))
    if rank == 0:
        print("ndb %9d world %d: identical=%s  nccl %.3f ms  peer %.3f ms (median of max over ranks)" % (
            ndb, world, bool(t.item()), res["nccl"], res["peer"]), flush=True)
if rank == 0:
    print("exchange_check", "OK" if ok else "MISMATCH", flush=True)
ex.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
