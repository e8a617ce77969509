"""N>1 host logic on CPU: two gloo ranks shard the database rows / the frames exactly as bench.py does on
the GPUs (c_orb_slam_b200.parallel), the per-rank partial results come from the CPU oracle (the checker),
and the gathered top-2 merged with the rule of merge_top2_kernel must equal the oracle on the whole set."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def merge_rule(idx, best, second):
    """Restatement (test side) of merge_top2_kernel: shards in ascending row order, the reference's update
    rule (src/ORBmatcher.cc:216-225) applied to (best, second) pairs."""
    world, nq = idx.shape
    o_i = np.full(nq, -1, np.int32); o_b = np.full(nq, 256, np.int32); o_s = np.full(nq, 256, np.int32)
    for k in range(world):
        for i in range(nq):
            if idx[k, i] < 0:
                continue
            if best[k, i] < o_b[i]:
                o_s[i] = min(o_b[i], second[k, i]); o_b[i] = best[k, i]; o_i[i] = idx[k, i]
            else:
                o_s[i] = min(o_s[i], best[k, i])
    return o_i, o_b, o_s


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from c_orb_slam_b200 import parallel, synth
    from oracle import pyoracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # --- database sharding + top-2 gather ---
        q, db = synth.make_match_set(96, 3001, seed=11)  # odd size: shards differ by one row
        lo, hi = parallel.shard_range(len(db), rank, world)
        i, b, s = O.match_bruteforce(q, db[lo:hi], threads=1)
        i = np.where(i >= 0, i + lo, -1).astype(np.int32)
        part = torch.from_numpy(np.stack([i, b, s]).astype(np.int32))
        g = parallel.gather_top2(part).numpy()  # [world][3][nq]
        # --- frame sharding: every rank extracts its block, counts are gathered (bench.py does the same) ---
        frames = synth.make_frames_fast(320, 240, 5, seed0=4242)
        flo, fhi = parallel.shard_range(len(frames), rank, world)
        ex = O.Extractor(300, 1.2, 4, 20, 7)
        mine = torch.zeros(len(frames), dtype=torch.int64)
        for f in range(flo, fhi):
            mine[f] = len(ex(frames[f])[0])
        dist.all_reduce(mine)
        if rank == 0:
            np.savez(os.path.join(out_dir, "r0.npz"), g=g, counts=mine.numpy())
    finally:
        dist.destroy_process_group()


def test_shard_range_properties():
    sys.path.insert(0, ROOT)
    from c_orb_slam_b200.parallel import shard_of, shard_range
    for n in (0, 1, 7, 8, 1024, 10_000_000):
        for world in (1, 2, 3, 4, 8):
            blocks = [shard_range(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[r][1] == blocks[r + 1][0] for r in range(world - 1))
            sizes = [hi - lo for lo, hi in blocks]
            assert max(sizes) - min(sizes) <= 1
    assert [shard_of(u, 10, 4) for u in range(10)] == [0, 0, 1, 1, 1, 2, 2, 3, 3, 3]
    with pytest.raises(ValueError):
        shard_range(10, 4, 4)


def test_two_rank_gloo_matches_single_process(tmp_path, oracle):
    import torch.multiprocessing as mp
    from c_orb_slam_b200 import synth
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r = np.load(tmp_path / "r0.npz")
    q, db = synth.make_match_set(96, 3001, seed=11)
    oi, ob, os_ = oracle.match_bruteforce(q, db, threads=2)
    mi, mb, ms = merge_rule(r["g"][:, 0], r["g"][:, 1], r["g"][:, 2])
    assert np.array_equal(mi, oi) and np.array_equal(mb, ob) and np.array_equal(ms, os_)
    frames = synth.make_frames_fast(320, 240, 5, seed0=4242)
    ex = oracle.Extractor(300, 1.2, 4, 20, 7)
    assert [len(ex(f)[0]) for f in frames] == r["counts"].tolist()
