"""Multi-GPU plumbing of the ORB front-end: one process per GPU, torch.distributed for the rendezvous.

Two paths shard:
  * extraction -- frames (and stereo pairs) are independent units: rank r takes the contiguous block
    shard_range(nframes, r, world); there is NO collective on the data path;
  * brute-force matching -- the descriptor database is sharded by rows; every rank scans its rows with
    the queries replicated, one all_gather of the per-rank (idx, best, second) [3][nq] int32 (12*nq bytes per
    rank) and the exact merge kernel (orb_gpu_merge_top2_device) give the global top-2.  Lowest index wins
    ties, as in the ascending scan of the reference's update rule (src/ORBmatcher.cc:216-225).
Windowed matching is "replicas only" (a call is microseconds of work).

The merge itself runs on the GPU (C ABI); nothing here computes descriptors or distances on the CPU.
"""
import torch
import torch.distributed as dist

from .api import lib, _check


def shard_range(n_units, rank, world):
    """Contiguous block [lo, hi) of `n_units` independent units (frames, stereo pairs, database rows) owned
    by `rank`: blocks are disjoint, cover [0, n_units) and differ in size by at most one."""
    if world < 1 or not (0 <= rank < world) or n_units < 0:
        raise ValueError("bad shard request: n_units=%r rank=%r world=%r" % (n_units, rank, world))
    return n_units * rank // world, n_units * (rank + 1) // world


def shard_of(unit, n_units, world):
    """Rank that owns `unit` under shard_range."""
    if not (0 <= unit < n_units):
        raise ValueError("unit out of range")
    r = min(world - 1, (unit * world + world - 1) // max(n_units, 1))
    while unit < shard_range(n_units, r, world)[0]:
        r -= 1
    while unit >= shard_range(n_units, r, world)[1]:
        r += 1
    return r


def gather_top2(part, group=None):
    """all_gather of the per-rank partial result part[3][nq] (idx, best, second; int32) -> [world][3][nq].
    Works on any backend (nccl on the GPUs, gloo in the CPU tests)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return part.unsqueeze(0)
    out = torch.empty(world * part.numel(), dtype=part.dtype, device=part.device)  # flat: accepted by nccl and gloo
    dist.all_gather_into_tensor(out, part.contiguous().view(-1), group=group)
    return out.view((world,) + tuple(part.shape))


def match_bruteforce_sharded(d_queries, d_db_shard, idx_base, group=None, stream=None):
    """Brute-force top-2 of the replicated queries against a row-sharded database.
    d_queries [nq,32] / d_db_shard [rows,32]: uint8 CUDA tensors; idx_base = global index of the shard's row 0.
    Returns a CUDA int32 tensor [3][nq] (global idx, best, second), identical on every rank."""
    if not d_queries.is_cuda or not d_db_shard.is_cuda:
        raise RuntimeError("match_bruteforce_sharded needs CUDA tensors: the product path has no CPU fallback")
    L = lib()
    nq = d_queries.shape[0]
    st = stream if stream is not None else torch.cuda.current_stream().cuda_stream
    part = torch.empty((3, nq), dtype=torch.int32, device=d_queries.device)
    _check(L.orb_gpu_match_bruteforce_device(d_queries.data_ptr(), nq, d_db_shard.data_ptr(), d_db_shard.shape[0],
                                             int(idx_base), part[0].data_ptr(), part[1].data_ptr(),
                                             part[2].data_ptr(), st), "orb_gpu_match_bruteforce_device")
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return part
    g = gather_top2(part, group).permute(1, 0, 2).contiguous()  # [3][world][nq], shards in ascending row order
    merged = torch.empty((3, nq), dtype=torch.int32, device=d_queries.device)
    _check(L.orb_gpu_merge_top2_device(g[0].data_ptr(), g[1].data_ptr(), g[2].data_ptr(), world, nq,
                                       merged[0].data_ptr(), merged[1].data_ptr(), merged[2].data_ptr(), st),
           "orb_gpu_merge_top2_device")
    return merged
