#!/usr/bin/env python
"""Under torchrun: aggregate pinned host->device copy rate when every rank copies at the same time (the bound of
the host-pointer path at N GPUs), with and without binding each process to the CPUs NVML reports as local to its GPU.
  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 tools/h2d_concurrent.py"""
import os, sys, time
import torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
NB = 157_286_400


def affinity_of_gpu(i):
    import pynvml as nv
    nv.nvmlInit()
    h = nv.nvmlDeviceGetHandleByIndex(i)
    words = nv.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
    cpus = [64 * w + b for w, v in enumerate(words) for b in range(64) if (v >> b) & 1]
    return cpus


def run(tag):
    h = torch.empty(NB, dtype=torch.uint8).pin_memory()
    h.fill_(1)
    d = torch.empty(NB, dtype=torch.uint8, device=dev)
    ho = torch.empty(31_459_328, dtype=torch.uint8).pin_memory()
    do = torch.empty(31_459_328, dtype=torch.uint8, device=dev)
    s2 = torch.cuda.Stream()
    for both in (False, True):
        for _ in range(2):
            d.copy_(h, non_blocking=True)
        torch.cuda.synchronize(); dist.barrier()
        t0 = time.perf_counter()
        for _ in range(10):
            d.copy_(h, non_blocking=True)
            if both:
                with torch.cuda.stream(s2):
                    ho.copy_(do, non_blocking=True)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], device=dev)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        if rank == 0:
            print("%s: %d ranks, H2D%s: %.2f ms per 157 MB copy -> %.1f GB/s per GPU, %.1f GB/s aggregate" % (
                tag, world, " + D2H 31 MB" if both else "", dt.item() / 10 * 1e3, NB * 10 / dt.item() / 1e9,
                world * NB * 10 / dt.item() / 1e9), flush=True)


if rank == 0:
    print("cpus visible:", sorted(os.sched_getaffinity(0)), flush=True)
run("unbound")
try:
    cpus = affinity_of_gpu(local)
    allowed = sorted(set(cpus) & os.sched_getaffinity(0))
    print("rank %d gpu %d local cpus %s -> usable %s" % (rank, local, cpus[:4] + ["..."] + cpus[-2:] if len(cpus) > 6 else cpus, allowed[:8]), flush=True)
    if allowed:
        os.sched_setaffinity(0, allowed)
        run("bound to the GPU's CPUs")
except Exception as e:
    print("affinity unavailable:", type(e).__name__, e, flush=True)
dist.destroy_process_group()
