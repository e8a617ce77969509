#!/usr/bin/env python
"""Per-iteration device time of the brute-force matcher (diagnostic for run-to-run variance)."""
import os, sys, time, subprocess, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import c_orb_slam_b200 as orb
from c_orb_slam_b200 import parallel
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(7)
db = torch.randint(0, 256, (10_000_000, 32), dtype=torch.uint8, device=dev, generator=g)
q = torch.randint(0, 256, (4096, 32), dtype=torch.uint8, device=dev, generator=g)
smi = []
def poll():
    while not stop:
        o = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.mem,power.draw,pstate,clocks_throttle_reasons.active",
                            "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
        smi.append(o)
        time.sleep(0.05)
stop = False
th = threading.Thread(target=poll); th.start()
ts = []
for i in range(12):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    parallel.match_bruteforce_sharded(q, db, 0)
    e1.record(); torch.cuda.synchronize()
    ts.append((round(e0.elapsed_time(e1), 2), round((time.perf_counter() - t0) * 1e3, 2)))
stop = True; th.join()
print("per-iteration (device ms, wall ms):", ts)
print("smi samples:", smi[::max(1, len(smi) // 8)])
print("popc peak", orb.popc_peak(0) / 1e9)
