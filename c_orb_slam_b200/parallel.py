"""Multi-GPU plumbing of the ORB front-end: one process per GPU, torch.distributed for the rendezvous.

Two paths shard:
  * extraction -- frames (and stereo pairs) are independent units: rank r takes the contiguous block
    shard_range(nframes, r, world); there is NO collective on the data path;
  * brute-force matching -- the descriptor database is sharded by rows; every rank scans its rows with
    the queries replicated; the per-rank (idx, best, second) [3][nq] int32 (12*nq bytes per rank) are exchanged
    and merged exactly.  Lowest index wins ties, as in the ascending scan of the reference's update rule
    (src/ORBmatcher.cc:216-225).  Two forms of the exchange: Top2Exchange -- ONE kernel per rank that stores
    its slice into every rank's gather buffer over NVLink peer memory, signals, waits and merges
    (orb_gpu_exchange_top2_device; no NCCL on the data path) -- and the plain one: NCCL all_gather +
    orb_gpu_merge_top2_device.
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


class Top2Exchange:
    """Peer-memory exchange of the per-rank top-2 (one per rank and process).  The constructor allocates this rank's
    gather buffer and arrival counters, all-gathers the CUDA IPC handles through torch.distributed and maps the
    peers' buffers; exchange() is then a single kernel launch per call."""

    def __init__(self, nq_max, device=None, group=None):
        import ctypes as C
        self._L = lib()
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.nq_max = int(nq_max)
        h = C.c_void_p()
        _check(self._L.orb_gpu_exchange_create(self.device, self.rank, self.world, self.nq_max, C.byref(h)),
               "orb_gpu_exchange_create")
        self._h = h
        if self.world > 1:
            nb = self._L.orb_gpu_exchange_handle_bytes()
            mine = torch.zeros(nb, dtype=torch.uint8)
            _check(self._L.orb_gpu_exchange_local_handle(self._h, mine.data_ptr()), "orb_gpu_exchange_local_handle")
            # the handles are 128 opaque bytes per rank: they go through the process group like any small tensor
            dev = torch.device("cuda", self.device) if dist.get_backend(group) == "nccl" else torch.device("cpu")
            allh = torch.empty(self.world * nb, dtype=torch.uint8, device=dev)
            dist.all_gather_into_tensor(allh, mine.to(dev), group=group)
            allh = allh.cpu().contiguous()
            _check(self._L.orb_gpu_exchange_connect(self._h, allh.data_ptr()), "orb_gpu_exchange_connect")
            dist.barrier(group)   # every rank has mapped every buffer before the first store

    def exchange(self, part, stream=None):
        """part: CUDA int32 [3][nq] (idx, best, second of this rank's rows) -> merged [3][nq], same on every rank."""
        nq = part.shape[1]
        out = torch.empty_like(part)
        st = stream if stream is not None else torch.cuda.current_stream().cuda_stream
        _check(self._L.orb_gpu_exchange_top2_device(self._h, part[0].data_ptr(), part[1].data_ptr(), part[2].data_ptr(),
                                                    nq, out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), st),
               "orb_gpu_exchange_top2_device")
        return out

    def match(self, d_queries, d_db_shard, idx_base, stream=None):
        """Search this rank's rows and exchange in ONE launch (orb_gpu_exchange_match_bruteforce_device): the block of
        the match kernel that finishes a query block stores its triples into every rank's gather buffer, signals, waits
        for the peers and merges.  -> global [3][nq], identical on every rank."""
        nq = d_queries.shape[0]
        out = torch.empty((3, nq), dtype=torch.int32, device=d_queries.device)
        st = stream if stream is not None else torch.cuda.current_stream().cuda_stream
        _check(self._L.orb_gpu_exchange_match_bruteforce_device(self._h, d_queries.data_ptr(), nq, d_db_shard.data_ptr(),
                                                                d_db_shard.shape[0], int(idx_base), out[0].data_ptr(),
                                                                out[1].data_ptr(), out[2].data_ptr(), st),
               "orb_gpu_exchange_match_bruteforce_device")
        return out

    def close(self):
        if getattr(self, "_h", None):
            self._L.orb_gpu_exchange_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def match_bruteforce_sharded(d_queries, d_db_shard, idx_base, group=None, stream=None, exchange=None, fused=True):
    """Brute-force top-2 of the replicated queries against a row-sharded database.
    d_queries [nq,32] / d_db_shard [rows,32]: uint8 CUDA tensors; idx_base = global index of the shard's row 0.
    exchange: a Top2Exchange (peer memory over NVLink: fused into the match kernel by default, fused=False runs the
    stand-alone exchange kernel after it) or None (NCCL all_gather + merge kernel).
    Returns a CUDA int32 tensor [3][nq] (global idx, best, second), identical on every rank."""
    if not d_queries.is_cuda or not d_db_shard.is_cuda:
        raise RuntimeError("match_bruteforce_sharded needs CUDA tensors: the product path has no CPU fallback")
    L = lib()
    nq = d_queries.shape[0]
    st = stream if stream is not None else torch.cuda.current_stream().cuda_stream
    if exchange is not None and fused and exchange.world > 1:
        return exchange.match(d_queries, d_db_shard, idx_base, st)   # search + exchange + merge: one kernel
    part = torch.empty((3, nq), dtype=torch.int32, device=d_queries.device)
    _check(L.orb_gpu_match_bruteforce_device(d_queries.data_ptr(), nq, d_db_shard.data_ptr(), d_db_shard.shape[0],
                                             int(idx_base), part[0].data_ptr(), part[1].data_ptr(),
                                             part[2].data_ptr(), st), "orb_gpu_match_bruteforce_device")
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return part
    if exchange is not None:
        return exchange.exchange(part, st)
    g = gather_top2(part, group).permute(1, 0, 2).contiguous()  # [3][world][nq], shards in ascending row order
    merged = torch.empty((3, nq), dtype=torch.int32, device=d_queries.device)
    _check(L.orb_gpu_merge_top2_device(g[0].data_ptr(), g[1].data_ptr(), g[2].data_ptr(), world, nq,
                                       merged[0].data_ptr(), merged[1].data_ptr(), merged[2].data_ptr(), st),
           "orb_gpu_merge_top2_device")
    return merged
