"""BASELINE.json's full sizes, checked through size-independent properties (the oracle would take minutes
there): config 5 -- 4096 queries x 10 M database rows; config 3 -- 1024 EuRoC frames in one batch."""
import numpy as np
import pytest

from c_orb_slam_b200 import synth

pytestmark = pytest.mark.gpu


def test_bruteforce_4096_x_10M_properties(gpu, oracle):
    import torch
    from c_orb_slam_b200 import parallel
    NQ, NDB = 4096, 10_000_000
    g = torch.Generator(device="cuda").manual_seed(5)
    d_db = torch.randint(0, 256, (NDB, 32), dtype=torch.uint8, device="cuda", generator=g)
    rng = np.random.default_rng(5)
    rows = rng.choice(NDB, NQ // 2, replace=False)
    flips = rng.integers(0, 31, NQ // 2)
    planted = synth.flip_bits(d_db[torch.from_numpy(rows).cuda()].cpu().numpy(), flips, rng)
    q = np.concatenate([planted, rng.integers(0, 256, (NQ - NQ // 2, 32), dtype=np.uint8)])
    d_q = torch.from_numpy(q).cuda()
    whole = parallel.match_bruteforce_sharded(d_q, d_db, 0).cpu().numpy()
    idx, best, second = whole
    # (1) planted rows are found with exactly the planted distance (random 256-bit rows sit ~128 bits away)
    assert np.array_equal(idx[:NQ // 2], rows) and np.array_equal(best[:NQ // 2], flips)
    # (2) order and range invariants of the update rule (src/ORBmatcher.cc:216-225)
    assert (best <= second).all() and (idx >= 0).all() and (idx < NDB).all() and (second <= 256).all()
    assert (second[:NQ // 2] > 60).all()
    # (3) the reported distance is the distance to the reported row (M1 on the gathered rows)
    got = d_db[torch.from_numpy(idx.astype(np.int64)).cuda()].cpu().numpy()
    assert np.array_equal(np.unpackbits(got ^ q, axis=1).sum(1), best)
    # (4) sharding invariance: 8 uneven row shards + exact merge == one pass (what 8 GPUs compute)
    L = gpu.lib()
    parts = []
    cuts = [0, 1_250_001, 2_500_000, 3_749_999, 5_000_000, 6_250_003, 7_500_000, 8_750_000, NDB]
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        parts.append(parallel.match_bruteforce_sharded(d_q, d_db[lo:hi], lo))
    gth = torch.stack(parts).permute(1, 0, 2).contiguous()
    out = torch.empty((3, NQ), dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert L.orb_gpu_merge_top2_device(gth[0].data_ptr(), gth[1].data_ptr(), gth[2].data_ptr(), 8, NQ,
                                       out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), st) == 0
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), whole)
    # (5) a slice the oracle finishes in seconds: 256 queries x the first 200 k rows
    oi, ob, os_ = oracle.match_bruteforce(q[:256], d_db[:200_000].cpu().numpy(), threads=8)
    sl = parallel.match_bruteforce_sharded(d_q[:256], d_db[:200_000], 0).cpu().numpy()
    assert np.array_equal(sl[0], oi) and np.array_equal(sl[1], ob) and np.array_equal(sl[2], os_)


def test_euroc_batch_1024_properties(gpu, oracle):
    """1024 EuRoC frames = 8 distinct frames x 128 copies, interleaved: every copy must give byte-identical
    keypoints and descriptors (no cross-frame interference in the batched kernels), and the 8 distinct results
    must equal the oracle."""
    import torch
    W, H = synth.SHAPES["euroc"]
    nfeat, B, D = 1200, 1024, 8
    cap = nfeat + 24
    base = synth.make_frames_fast(W, H, D, seed0=900)
    d_base = torch.from_numpy(base).cuda()
    d_fr = d_base[torch.arange(B, device="cuda") % D].contiguous()
    ex = gpu.ORBextractor(nfeat, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
    d_k = torch.zeros((B, cap, 28), dtype=torch.uint8, device="cuda")
    d_d = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(B, dtype=torch.int32, device="cuda")
    d_s = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    ex.extract_batch_device(d_fr.data_ptr(), B, W, H, W, W * H, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(),
                            d_s.data_ptr(), st)
    torch.cuda.synchronize()
    assert int(d_s.item()) == 0
    n = d_n.cpu().numpy()
    assert (n.reshape(-1, D) == n[:D]).all() and (n >= nfeat).all()
    k = d_k.view(B // D, D, cap, 28)
    d = d_d.view(B // D, D, cap, 32)
    for j in range(D):
        nj = int(n[j])
        assert bool((k[:, j, :nj] == k[0, j, :nj]).all()) and bool((d[:, j, :nj] == d[0, j, :nj]).all())
    ex_o = oracle.Extractor(nfeat, 1.2, 8, 20, 7)
    for j in range(D):
        ko, do = ex_o(base[j])
        kg = d_k[j, :len(ko)].cpu().numpy().view(gpu.KP_DTYPE).reshape(-1)
        assert len(ko) == n[j]
        for f in ("x", "y", "size", "response", "octave", "angle"):
            assert np.array_equal(ko[f], kg[f]), f
        assert np.array_equal(do, d_d[j, :len(ko)].cpu().numpy())


def test_4k_frame_with_more_than_32768_candidates_per_level(gpu, oracle):
    """A 3840 x 2160 textured frame has ~57 k FAST candidates on level 0 -- more than the 4096 keys the quadtree sorts
    in shared memory and more than the 32768 its arrays held in round 1: the global-memory key path (bitonic sort)
    gives the reference's keypoints and descriptors (src/ORBextractor.cc:539-763 on a large level)."""
    W, H = 3840, 2160
    img = synth.make_frame(W, H, 5)
    ex = gpu.ORBextractor(5000, 1.2, 8, 20, 7, max_width=W, max_height=H)
    k, d = ex(img)
    ex_o = oracle.Extractor(5000, 1.2, 8, 20, 7)
    ko, do = ex_o(img)
    assert len(ex_o.candidates(0)) > 32768
    assert len(k) == len(ko) and np.array_equal(k, ko) and np.array_equal(d, do)
