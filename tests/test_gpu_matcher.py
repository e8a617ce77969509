"""GPU parity of the Hamming matchers against the CPU oracle (pytest -m gpu): bit-exact."""
import numpy as np
import pytest

from c_orb_slam_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nq,ndb", [(1, 1), (7, 300), (512, 20000), (1000, 70001)])
@pytest.mark.parametrize("variant", [(0, 1), (1, 2), (1, 4), (2, 2), (3, 2)])  # 8-POPC, carry-save x2, mma.sync, tcgen05
def test_bruteforce_top2(gpu, oracle, nq, ndb, variant):
    q, db = synth.make_match_set(nq, ndb, seed=11)
    gpu.lib().orb_gpu_set_bruteforce_variant(*variant)
    m = gpu.ORBmatcher(0.9)
    idx, best, second = m.match_bruteforce(q, db)
    oi, ob, os_ = oracle.match_bruteforce(q, db, threads=8)
    assert np.array_equal(best, ob) and np.array_equal(second, os_) and np.array_equal(idx, oi)
    gpu.lib().orb_gpu_set_bruteforce_variant(3, 2)   # back to the default (tcgen05 form)
    # the accept flags come from the device kernel (orb_gpu_ratio_test_device), checked against the float rule
    acc = m.accept_device(best, second, th=m.TH_HIGH)
    ref = (ob <= 100) & (ob.astype(np.float32) < np.float32(0.9) * os_.astype(np.float32))
    assert np.array_equal(acc, ref)


def test_hamming_pairs_device_equals_descriptor_distance(gpu, oracle):
    """M1: orb_gpu_hamming_pairs_device against the reference's own ORBmatcher::DescriptorDistance
    (src/ORBmatcher.cc:1647-1663, verbatim build) on 10^5 random pairs plus the all-0 / all-1 extremes."""
    import torch
    rng = np.random.default_rng(77)
    n = 100000
    a = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    b = synth.flip_bits(a, rng.integers(0, 257, n), rng)
    b[n // 2:] = rng.integers(0, 256, (n - n // 2, 32), dtype=np.uint8)
    a[0] = 0; b[0] = 0; a[1] = 0; b[1] = 255; a[2] = 255; b[2] = 255; a[3] = 255; b[3] = 0
    da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    out = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert gpu.lib().orb_gpu_hamming_pairs_device(da.data_ptr(), db.data_ptr(), n, out.data_ptr(), st) == 0
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    want = np.unpackbits(a ^ b, axis=1).sum(1).astype(np.int32)
    assert np.array_equal(got, want) and got[1] == 256 and got[3] == 256 and got[0] == 0
    impl = "reference" if oracle.have_ref_matcher() else "oracle"
    for i in list(range(8)) + rng.integers(0, n, 2000).tolist():
        assert oracle.descriptor_distance(a[i], b[i], impl=impl) == got[i]
    assert gpu.lib().orb_gpu_hamming_pairs_device(da.data_ptr(), db.data_ptr(), 0, out.data_ptr(), st) == 0  # empty


@pytest.mark.parametrize("nnratio", [0.6, 0.75, 0.9])
@pytest.mark.parametrize("th", [50, 100])
def test_ratio_test_device_all_pairs(gpu, nnratio, th):
    """orb_gpu_ratio_test_device on all 257 x 257 (best, second) pairs against the float rule of
    src/ORBmatcher.cc:228-230: best <= th && (float)best < mfNNratio * (float)second."""
    best, second = [x.ravel().astype(np.int32) for x in np.meshgrid(np.arange(257), np.arange(257), indexing="ij")]
    acc = gpu.ORBmatcher(nnratio).accept_device(best, second, th=th)
    want = (best <= th) & (best.astype(np.float32) < np.float32(nnratio) * second.astype(np.float32))
    assert np.array_equal(acc, want) and want.sum() > 1000


@pytest.mark.parametrize("nq,ndb", [(1, 255), (255, 257), (256, 148 * 256 - 1), (257, 148 * 256 + 1), (300, 148 * 768),
                                    (513, 49 * 25 * 256), (513, 49 * 25 * 256 + 1), (100, 300001)])
def test_bruteforce_tcgen05_ragged_slices(gpu, oracle, nq, ndb):
    """The tcgen05 kernel's MMA issuer walks the tiles of a slice two at a time (one per B stage): slices of one, an odd
    and an even number of 256-row tiles, last tiles of a single row, query blocks with one live query."""
    q, db = synth.make_match_set(nq, ndb, seed=nq + ndb)
    gpu.lib().orb_gpu_set_bruteforce_variant(3, 2)
    m = gpu.ORBmatcher(0.9)
    for rep in range(2):  # the second call reuses the slice tickets the first one reset
        idx, best, second = m.match_bruteforce(q, db)
        oi, ob, os_ = oracle.match_bruteforce(q, db, threads=8)
        assert np.array_equal(best, ob) and np.array_equal(second, os_) and np.array_equal(idx, oi)


@pytest.mark.parametrize("variant", [(2, 2), (3, 2)])
def test_bruteforce_extreme_descriptors(gpu, oracle, variant):
    """All-zero / all-one descriptors (S = +-256, the ends of the accumulator range), ragged sizes, planted ties."""
    rng = np.random.default_rng(5)
    q = synth.random_descriptors(259, 3)
    q[0] = 0; q[1] = 255; q[2] = 0x0f; q[3] = 0xf0
    db = synth.random_descriptors(1031, 4)
    db[5] = 0; db[6] = 255; db[700] = 255; db[701] = 0; db[1030] = q[258]; db[9] = q[258]
    db[300] = q[7]; db[301] = q[7]; db[302] = q[7] ^ np.uint8(1)
    gpu.lib().orb_gpu_set_bruteforce_variant(*variant)
    try:
        idx, best, second = gpu.ORBmatcher(0.9).match_bruteforce(q, db)
    finally:
        gpu.lib().orb_gpu_set_bruteforce_variant(3, 2)
    oi, ob, os_ = oracle.match_bruteforce(q, db, threads=4)
    assert np.array_equal(best, ob) and np.array_equal(second, os_) and np.array_equal(idx, oi)


def test_bruteforce_empty_and_ties(gpu, oracle):
    m = gpu.ORBmatcher()
    q = synth.random_descriptors(5, 1)
    idx, best, second = m.match_bruteforce(q, np.empty((0, 32), np.uint8))
    assert (idx == -1).all() and (best == 256).all() and (second == 256).all()
    db = np.repeat(q[:1], 9, axis=0)                      # nine identical rows: lowest index wins
    idx, best, second = m.match_bruteforce(q[:1], db)
    assert idx[0] == 0 and best[0] == 0 and second[0] == 0
    inv = (~q[:1]).astype(np.uint8)                       # distance 256 never beats the initial 256
    idx, best, second = m.match_bruteforce(q[:1], inv)
    oi, ob, os_ = oracle.match_bruteforce(q[:1], inv)
    assert idx[0] == oi[0] == -1 and best[0] == ob[0] == 256


def _frame_from_extractor(oracle, W, H, seed, nfeat):
    ex = oracle.Extractor(nfeat, 1.2, 8, 20, 7)
    kps, desc = ex(synth.make_frame(W, H, seed))
    return ex, kps, desc


def _projection_case(oracle, seed, th, stereo):
    W, H = 640, 480
    ex, kps, desc = _frame_from_extractor(oracle, W, H, seed, 2000)
    rng = np.random.default_rng(seed)
    nk = len(kps)
    sf = ex.tables()["scale"]
    M = 2000
    proj = np.zeros(M, oracle.PROJ_DTYPE)
    qdesc = np.empty((M, 32), np.uint8)
    src = rng.integers(0, nk, M)
    near = rng.random(M) < 0.75
    proj["u"] = np.where(near, kps["x"][src] + rng.normal(0, 2, M), rng.uniform(0, W, M)).astype(np.float32)
    proj["v"] = np.where(near, kps["y"][src] + rng.normal(0, 2, M), rng.uniform(0, H, M)).astype(np.float32)
    proj["level"] = np.clip(kps["octave"][src] + rng.integers(0, 2, M), 0, 7)
    proj["view_cos"] = rng.uniform(0.99, 1.0, M).astype(np.float32)
    proj["valid"] = rng.random(M) < 0.95
    proj["has_obs"] = rng.random(M) < 0.9
    qdesc[:] = synth.flip_bits(desc[src], rng.integers(0, 61, M), rng)
    # duplicates: several map points aimed at the same keypoint -> ordered claims matter
    dup = rng.integers(0, M, 300)
    proj[dup] = proj[(dup + 1) % M]
    qdesc[dup] = qdesc[(dup + 1) % M]
    u_right = None
    if stereo:
        u_right = np.where(rng.random(nk) < 0.6, kps["x"] - rng.uniform(1, 40, nk), -1).astype(np.float32)
        proj["ur"] = (proj["u"] - rng.uniform(1, 40, M)).astype(np.float32)
    blocked = (rng.random(nk) < 0.05).astype(np.uint8)
    grid = (0.0, 0.0, np.float32(64) / np.float32(W), np.float32(48) / np.float32(H))
    return kps, desc, proj, qdesc, u_right, blocked, grid, sf


@pytest.mark.parametrize("th,stereo,seed", [(1.0, False, 1), (3.0, False, 2), (5.0, True, 3), (15.0, True, 4)])
def test_search_by_projection(gpu, oracle, th, stereo, seed):
    kps, desc, proj, qdesc, u_right, blocked, grid, sf = _projection_case(oracle, seed, th, stereo)
    n_o, match_o, blocked_o = oracle.search_by_projection(proj, qdesc, kps, desc, u_right, grid, sf, th, 0.9,
                                                          100, blocked.copy())
    frame = dict(kps=kps, desc=desc, u_right=u_right, grid=grid, scale_factors=sf, blocked=blocked.copy())
    pg = np.zeros(len(proj), gpu.PROJ_DTYPE)
    for f in ("u", "v", "ur", "level", "view_cos", "valid", "has_obs"):
        pg[f] = proj[f]
    n_g, match_g = gpu.ORBmatcher(0.9).SearchByProjection(frame, dict(proj=pg, desc=qdesc), th)
    assert n_o > 100
    assert n_g == n_o and np.array_equal(match_g, match_o) and np.array_equal(frame["blocked"], blocked_o)


@pytest.mark.parametrize("th,stereo,seed", [(1.0, False, 1), (5.0, True, 3)])
def test_search_by_projection_resident_frame(gpu, oracle, th, stereo, seed):
    """The same search against a frame kept in device memory (orb_gpu_frame_*): uploaded once, searched three times
    with different blocked flags -- results identical to the oracle and to the host-pointer call."""
    kps, desc, proj, qdesc, u_right, blocked, grid, sf = _projection_case(oracle, seed, th, stereo)
    pg = np.zeros(len(proj), gpu.PROJ_DTYPE)
    for f in ("u", "v", "ur", "level", "view_cos", "valid", "has_obs"):
        pg[f] = proj[f]
    res = gpu.GpuFrame().upload(kps, desc, u_right, grid)
    assert len(res) == len(kps)
    rng = np.random.default_rng(seed)
    for rep in range(3):
        b0 = blocked if rep == 0 else (rng.random(len(kps)) < 0.1 * rep).astype(np.uint8)
        n_o, match_o, blocked_o = oracle.search_by_projection(proj, qdesc, kps, desc, u_right, grid, sf, th, 0.9, 100,
                                                              b0.copy())
        frame = dict(resident=res, scale_factors=sf, blocked=b0.copy())
        n_g, match_g = gpu.ORBmatcher(0.9).SearchByProjection(frame, dict(proj=pg, desc=qdesc), th)
        assert n_g == n_o and np.array_equal(match_g, match_o) and np.array_equal(frame["blocked"], blocked_o)


@pytest.mark.parametrize("mode,th_accept,check_ori,seed", [
    ("last_forward", 100, True, 2), ("kf", 64, True, 5), ("sim3", 50, False, 6)])
def test_search_windowed_resident_frame(gpu, oracle, mode, th_accept, check_ori, seed):
    kps, desc, q, qdesc, u_right, blocked, grid = _window_case(oracle, seed, mode)
    gate = 1 if u_right is not None else 0
    n_o, match_o, blocked_o, mq_o = oracle.search_windowed(q, qdesc, kps, desc, u_right, grid, th_accept, gate,
                                                           check_ori, blocked.copy())
    res = gpu.GpuFrame().upload(kps, desc, u_right, grid)
    for _ in range(2):
        frame = dict(resident=res, blocked=blocked.copy())
        n_g, match_g, mq_g = gpu.ORBmatcher(0.9, checkOri=check_ori).SearchWindowed(
            frame, _to_gpu_queries(gpu, q), qdesc, th_accept, gate)
        assert n_g == n_o and np.array_equal(match_g, match_o) and np.array_equal(frame["blocked"], blocked_o)
        assert np.array_equal(mq_g, mq_o)


def test_resident_frame_from_extractor_outputs(gpu, oracle):
    """Keypoints and descriptors go from orb_gpu_extract_batch_device straight into a resident frame (device to
    device); a search against it equals the search against the same arrays brought through the host."""
    import torch
    W, H, nfeat = 640, 480, 2000
    cap = nfeat + 24
    img = synth.make_frame(W, H, 77)
    ex = gpu.ORBextractor(nfeat, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=2)
    d_fr = torch.from_numpy(np.stack([img, img[::-1].copy()])).cuda()
    d_k = torch.empty((2, cap, 28), dtype=torch.uint8, device="cuda")
    d_d = torch.empty((2, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(2, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    ex.extract_batch_device(d_fr.data_ptr(), 2, W, H, W, W * H, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), None, st)
    torch.cuda.synchronize()   # a NULL stream means the handle's own (non-blocking) stream: the call is asynchronous
    n = d_n.cpu().numpy()
    grid = (0.0, 0.0, np.float32(64) / np.float32(W), np.float32(48) / np.float32(H))
    f = 1
    res = gpu.GpuFrame().from_device(d_k[f].data_ptr(), d_d[f].data_ptr(), None, int(n[f]), grid, st)
    kps = d_k[f, :n[f]].cpu().numpy().view(gpu.KP_DTYPE).reshape(-1)
    desc = d_d[f, :n[f]].cpu().numpy()
    rng = np.random.default_rng(9)
    M = 1500
    src = rng.integers(0, len(kps), M)
    q = np.zeros(M, oracle.WINDOW_DTYPE)
    q["u"] = (kps["x"][src] + rng.normal(0, 3, M)).astype(np.float32)
    q["v"] = (kps["y"][src] + rng.normal(0, 3, M)).astype(np.float32)
    q["radius"] = 12.0
    q["min_level"], q["max_level"] = kps["octave"][src] - 1, kps["octave"][src] + 1
    q["angle"] = kps["angle"][src]
    q["valid"] = 1
    q["blocks"] = 1
    qdesc = synth.flip_bits(desc[src], rng.integers(0, 50, M), rng)
    m = gpu.ORBmatcher(0.9, checkOri=True)
    qg = _to_gpu_queries(gpu, q)
    n_a, match_a, mq_a = m.SearchWindowed(dict(resident=res), qg, qdesc, 100, 0)
    n_b, match_b, mq_b = m.SearchWindowed(dict(kps=kps, desc=desc, u_right=None, grid=grid), qg, qdesc, 100, 0)
    n_o, match_o, _, mq_o = oracle.search_windowed(q, qdesc, kps, desc, None, grid, 100, 0, True, np.zeros(len(kps), np.uint8))
    assert n_a > 500 and n_a == n_b == n_o and np.array_equal(match_a, match_b) and np.array_equal(match_a, match_o)
    assert np.array_equal(mq_a, mq_o)


def test_stereo_match(gpu, oracle):
    W, H = synth.SHAPES["kitti"]
    left, right = synth.make_stereo_pair(W, H, 31)
    exl = oracle.Extractor(2000, 1.2, 8, 20, 7)
    kl, dl = exl(left)
    kr, dr = oracle.Extractor(2000, 1.2, 8, 20, 7)(right)
    sf = exl.tables()["scale"]
    max_d = np.float32(718.856)  # bf/mb = fx (KITTI00-02.yaml; Frame.cc:496-498)
    bi_o, bd_o = oracle.stereo_match(kl, dl, kr, dr, H, sf, 0.0, max_d)
    bi_g, bd_g = gpu.ORBmatcher().stereo_match(kl, dl, kr, dr, H, sf, 0.0, max_d)
    assert (bi_o >= 0).sum() > 200
    assert np.array_equal(bi_g, bi_o) and np.array_equal(bd_g, bd_o)


def test_sharded_bruteforce_merge_on_device(gpu, oracle):
    """Database split into 3 row shards on one GPU: per-shard device match + merge kernel == whole database
    (the N>1 path of c_orb_slam_b200.parallel without the all_gather; gloo covers that on CPU)."""
    import torch
    from c_orb_slam_b200 import parallel
    q, db = synth.make_match_set(300, 25003, seed=3)
    dq = torch.from_numpy(q).cuda()
    L = gpu.lib()
    parts = []
    for r in range(3):
        lo, hi = parallel.shard_range(len(db), r, 3)
        parts.append(parallel.match_bruteforce_sharded(dq, torch.from_numpy(db[lo:hi]).cuda(), lo))
    g = torch.stack(parts).permute(1, 0, 2).contiguous()  # [3][shards][nq]
    out = torch.empty((3, len(q)), dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert L.orb_gpu_merge_top2_device(g[0].data_ptr(), g[1].data_ptr(), g[2].data_ptr(), 3, len(q),
                                       out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), st) == 0
    torch.cuda.synchronize()
    oi, ob, os_ = oracle.match_bruteforce(q, db, threads=8)
    o = out.cpu().numpy()
    assert np.array_equal(o[0], oi) and np.array_equal(o[1], ob) and np.array_equal(o[2], os_)


def _window_case(oracle, seed, mode):
    """Queries of the best-only projection searches as their callers build them (M3: last-frame points,
    forward / backward / neutral level windows, stereo gate; M4/M5: level +-1 / [l-1, l], no gate)."""
    W, H = 640, 480
    ex, kps, desc = _frame_from_extractor(oracle, W, H, seed, 2000)
    rng = np.random.default_rng(100 + seed)
    nk, M = len(kps), 1800
    sf = ex.tables()["scale"]
    src = rng.integers(0, nk, M)
    q = np.zeros(M, oracle.WINDOW_DTYPE)
    near = rng.random(M) < 0.8
    q["u"] = np.where(near, kps["x"][src] + rng.normal(0, 3, M), rng.uniform(0, W, M)).astype(np.float32)
    q["v"] = np.where(near, kps["y"][src] + rng.normal(0, 3, M), rng.uniform(0, H, M)).astype(np.float32)
    lvl = kps["octave"][src]
    th = {"last_mono": 15.0, "last_forward": 7.0, "last_backward": 7.0, "kf": 10.0, "sim3": 10.0}[mode]
    q["radius"] = (np.float32(th) * sf[lvl]).astype(np.float32)
    if mode == "last_forward":
        q["min_level"], q["max_level"] = lvl, -1
    elif mode == "last_backward":
        q["min_level"], q["max_level"] = 0, lvl
    elif mode == "sim3":
        q["min_level"], q["max_level"] = lvl - 1, lvl
    else:
        q["min_level"], q["max_level"] = lvl - 1, lvl + 1
    # source angles: mostly a common rotation of ~40 degrees, some outliers -> the histogram prunes
    rot = np.where(rng.random(M) < 0.85, rng.normal(40, 6, M), rng.uniform(0, 360, M))
    q["angle"] = np.mod(kps["angle"][src] + rot, 360).astype(np.float32)
    q["valid"] = rng.random(M) < 0.95
    q["blocks"] = (rng.random(M) < 0.9) if mode.startswith("last") else 1
    qdesc = synth.flip_bits(desc[src], rng.integers(0, 70, M), rng)
    dup = rng.integers(0, M, 300)   # several points aimed at one keypoint: the claim order matters
    q[dup] = q[(dup + 1) % M]
    qdesc[dup] = qdesc[(dup + 1) % M]
    u_right = None
    if mode in ("last_forward", "last_backward"):
        u_right = np.where(rng.random(nk) < 0.6, kps["x"] - rng.uniform(1, 40, nk), -1).astype(np.float32)
        q["ur"] = (q["u"] - rng.uniform(1, 40, M)).astype(np.float32)
    blocked = (rng.random(nk) < 0.05).astype(np.uint8)
    grid = (0.0, 0.0, np.float32(64) / np.float32(W), np.float32(48) / np.float32(H))
    return kps, desc, q, qdesc, u_right, blocked, grid


def _to_gpu_queries(gpu, q):
    qg = np.zeros(len(q), gpu.WINDOW_DTYPE)
    for f in ("u", "v", "radius", "ur", "min_level", "max_level", "angle", "valid", "blocks"):
        qg[f] = q[f]
    return qg


@pytest.mark.parametrize("mode,th_accept,check_ori,seed", [
    ("last_mono", 100, True, 1), ("last_forward", 100, True, 2), ("last_backward", 100, True, 3),
    ("kf", 100, True, 4), ("kf", 64, True, 5), ("sim3", 50, False, 6)])
def test_search_windowed_variants(gpu, oracle, mode, th_accept, check_ori, seed):
    """M3 / M4 / M5 + M8: best-only windowed search, ordered claims, rotation histogram -- bit-exact."""
    kps, desc, q, qdesc, u_right, blocked, grid = _window_case(oracle, seed, mode)
    gate = 1 if u_right is not None else 0
    n_o, match_o, blocked_o, mq_o = oracle.search_windowed(q, qdesc, kps, desc, u_right, grid, th_accept, gate,
                                                           check_ori, blocked.copy())
    frame = dict(kps=kps, desc=desc, u_right=u_right, grid=grid, blocked=blocked.copy())
    n_g, match_g, mq_g = gpu.ORBmatcher(0.9, checkOri=check_ori).SearchWindowed(
        frame, _to_gpu_queries(gpu, q), qdesc, th_accept, gate)
    assert n_o > 100
    assert n_g == n_o and np.array_equal(match_g, match_o) and np.array_equal(frame["blocked"], blocked_o)
    assert np.array_equal(mq_g, mq_o)
    # consistency of the two views: a surviving claim points back at its query
    won = np.flatnonzero(match_g >= 0)
    assert np.array_equal(mq_g[match_g[won]], won)


@pytest.mark.parametrize("stereo,seed", [(False, 7), (True, 8)])
def test_fuse_search_chi_square_gate(gpu, oracle, stereo, seed):
    """Search loop of ORBmatcher::Fuse (src/ORBmatcher.cc:825-975): level window [l-1, l], chi-square gate
    5.99 / 7.8 on the reprojection error, best <= TH_LOW, no claims -- one keypoint may be chosen by several
    map points (the caller then Replace()s / AddMapPoint()s in query order)."""
    kps, desc, q, qdesc, _, blocked, grid = _window_case(oracle, seed, "sim3")
    rng = np.random.default_rng(seed)
    nk, M = len(kps), len(q)
    q["blocks"] = 0
    q["radius"] = (q["radius"] * np.float32(0.3)).astype(np.float32)   # th = 3 in LocalMapping's fuse
    u_right = None
    if stereo:
        u_right = np.where(rng.random(nk) < 0.6, kps["x"] - rng.uniform(1, 40, nk), -1).astype(np.float32)
        near = rng.random(M) < 0.5
        q["ur"] = np.where(near, q["u"] - 20, q["u"] - rng.uniform(1, 40, M)).astype(np.float32)
    sigma2 = (np.float32(1.2) ** np.arange(8, dtype=np.float32)) ** 2
    inv_sigma2 = (np.float32(1.0) / sigma2).astype(np.float32)
    blocked[:] = 0
    n_o, match_o, blocked_o, mq_o = oracle.search_windowed(q, qdesc, kps, desc, u_right, grid, 50, 2, False,
                                                           blocked.copy(), inv_sigma2)
    frame = dict(kps=kps, desc=desc, u_right=u_right, grid=grid, blocked=blocked.copy())
    n_g, match_g, mq_g = gpu.ORBmatcher(0.9, checkOri=False).SearchWindowed(
        frame, _to_gpu_queries(gpu, q), qdesc, 50, 2, inv_sigma2)
    assert (mq_o >= 0).sum() > 100
    # the gate must bite: the same search without it accepts strictly more
    _, _, _, mq_nogate = oracle.search_windowed(q, qdesc, kps, desc, u_right, grid, 50, 0, False, blocked.copy())
    assert (mq_nogate >= 0).sum() > (mq_o >= 0).sum()
    assert n_g == n_o and np.array_equal(mq_g, mq_o) and np.array_equal(match_g, match_o)
    assert not frame["blocked"].any() and not blocked_o.any()


def test_search_by_sim3_two_passes_agree(gpu, oracle):
    """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1102-1326): two claim-free windowed passes (KF1 points into
    KF2 and back, level window [l-1, l], best <= TH_HIGH) and the mutual check vnMatch2[vnMatch1[i]] == i."""
    kps1, desc1, q12, qd12, _, _, grid = _window_case(oracle, 9, "sim3")
    kps2, desc2, q21, qd21, _, _, _ = _window_case(oracle, 10, "sim3")
    for q in (q12, q21):
        q["blocks"] = 0
    m = gpu.ORBmatcher(0.9, checkOri=False)
    res = []
    for kps, desc, q, qd in ((kps2, desc2, q12, qd12), (kps1, desc1, q21, qd21)):
        z = np.zeros(len(kps), np.uint8)
        _, _, _, mq_o = oracle.search_windowed(q, qd, kps, desc, None, grid, 100, 0, False, z.copy())
        frame = dict(kps=kps, desc=desc, u_right=None, grid=grid, blocked=z.copy())
        _, _, mq_g = m.SearchWindowed(frame, _to_gpu_queries(gpu, q), qd, 100, 0)
        assert np.array_equal(mq_g, mq_o) and (mq_g >= 0).sum() > 100
        res.append(mq_g)
    # the mutual check is host bookkeeping over the two result vectors; exercised here on index vectors
    n = min(len(res[0]), len(res[1]))
    a, b = res[0][:n] % n, res[1][:n] % n
    mutual = np.flatnonzero((res[0][:n] >= 0) & (b[a] == np.arange(n)))
    assert mutual.size >= 0


def test_features_in_area_order(gpu, oracle):
    """N2 / M11: the device grid returns every window's candidates in the order of Frame::GetFeaturesInArea
    (src/Frame.cc:327-380) -- the order the stable top-2 tie-breaks depend on."""
    W, H = 640, 480
    _, kps, _ = _frame_from_extractor(oracle, W, H, 21, 2000)
    rng = np.random.default_rng(21)
    grid = (0.0, 0.0, np.float32(64) / np.float32(W), np.float32(48) / np.float32(H))
    M = 400
    q = np.zeros(M, gpu.WINDOW_DTYPE)
    q["u"] = rng.uniform(-30, W + 30, M).astype(np.float32)
    q["v"] = rng.uniform(-30, H + 30, M).astype(np.float32)
    q["radius"] = rng.choice([3.0, 10.0, 40.0, 100.0, 700.0], M).astype(np.float32)
    lv = rng.integers(0, 8, M)
    mode = rng.integers(0, 4, M)
    q["min_level"] = np.where(mode == 0, -1, np.where(mode == 1, lv, np.where(mode == 2, 0, lv - 1)))
    q["max_level"] = np.where(mode == 0, -1, np.where(mode == 1, -1, np.where(mode == 2, lv, lv + 1)))
    q["valid"] = 1
    got = gpu.ORBmatcher().GetFeaturesInArea(dict(kps=kps, grid=grid), q)
    total = 0
    for i in range(M):
        want = oracle.features_in_area(kps, grid, q["u"][i], q["v"][i], q["radius"][i], int(q["min_level"][i]),
                                       int(q["max_level"][i]))
        assert np.array_equal(got[i], want), i
        total += len(want)
    assert total > 5000 and any(len(g) == 0 for g in got)


def test_peer_memory_exchange_many_queries(gpu):
    """65536 queries per call: more query blocks than could ever be resident next to each other with one block per
    256 queries -- the exchange kernel caps its grid and strides (every block waits for all blocks of all ranks)."""
    import ctypes as C
    import torch
    L = gpu.lib()
    nq, world = 65536, 2
    hs = (C.c_void_p * world)()
    for r in range(world):
        h = C.c_void_p()
        assert L.orb_gpu_exchange_create(0, r, world, nq, C.byref(h)) == 0
        hs[r] = h
    assert L.orb_gpu_exchange_connect_local(hs, world) == 0
    streams = [torch.cuda.Stream() for _ in range(world)]
    rng = np.random.default_rng(3)
    try:
        for call in range(3):
            n = nq - 777 * call
            parts = []
            for r in range(world):
                best = rng.integers(0, 120, n).astype(np.int32)
                second = (best + rng.integers(0, 40, n)).astype(np.int32)
                idx = (r * 1000000 + rng.integers(0, 1000000, n)).astype(np.int32)
                none = rng.random(n) < 0.01
                idx[none] = -1; best[none] = 256; second[none] = 256
                parts.append(np.stack([idx, best, second]))
            b0, b1 = parts[0][1], parts[1][1]
            first = b0 <= b1                       # rank 0 holds the lower rows: it wins ties
            want = np.where(first, parts[0], parts[1])
            want[2] = np.where(first, np.minimum(parts[0][2], b1), np.minimum(parts[1][2], b0))
            outs = []
            dparts = [torch.from_numpy(p).cuda() for p in parts]
            torch.cuda.synchronize()
            for r in range(world):
                out = torch.empty((3, n), dtype=torch.int32, device="cuda")
                p = dparts[r]
                assert L.orb_gpu_exchange_top2_device(hs[r], p[0].data_ptr(), p[1].data_ptr(), p[2].data_ptr(), n,
                                                      out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(),
                                                      streams[r].cuda_stream) == 0
                outs.append(out)
            torch.cuda.synchronize()
            for out in outs:
                assert np.array_equal(out.cpu().numpy(), want)
    finally:
        torch.cuda.synchronize()
        for r in range(world):
            L.orb_gpu_exchange_destroy(hs[r])


@pytest.mark.parametrize("nq,ndb", [(1000, 60000), (256, 9000), (4096, 300000)])
def test_fused_match_and_exchange_two_ranks_on_one_gpu(gpu, oracle, nq, ndb):
    """orb_gpu_exchange_match_bruteforce_device: search + exchange + merge in ONE launch per rank.  Both ranks of a
    world of 2 live in this process on one GPU (two streams); the database is cut at uneven places; five calls in a
    row reuse the double-buffered gather slots and the per-query-block sequence numbers, the fourth is a plain
    exchange call in between (the two forms may be mixed).  Every rank must hold the oracle's global top-2."""
    import ctypes as C
    import torch
    L = gpu.lib()
    before = gpu.launch_count()
    hs = (C.c_void_p * 2)()
    for r in range(2):
        h = C.c_void_p()
        assert L.orb_gpu_exchange_create(0, r, 2, 4096, C.byref(h)) == 0
        hs[r] = h
    assert L.orb_gpu_exchange_connect_local(hs, 2) == 0
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    try:
        for call in range(5):
            q, db = synth.make_match_set(nq, ndb, seed=70 + call)
            want = np.stack(oracle.match_bruteforce(q, db, threads=8))
            dq, ddb = torch.from_numpy(q).cuda(), torch.from_numpy(db).cuda()
            cut = ndb // 3 + 23 * call
            torch.cuda.synchronize()
            outs = []
            if call == 3:   # the stand-alone exchange kernel between fused calls
                parts = []
                for lo, hi in ((0, cut), (cut, ndb)):
                    p = torch.empty((3, nq), dtype=torch.int32, device="cuda")
                    assert L.orb_gpu_match_bruteforce_device(dq.data_ptr(), nq, ddb[lo:hi].data_ptr(), hi - lo, lo,
                                                             p[0].data_ptr(), p[1].data_ptr(), p[2].data_ptr(), None) == 0
                    parts.append(p)
                torch.cuda.synchronize()
                for r in range(2):
                    out = torch.empty((3, nq), dtype=torch.int32, device="cuda")
                    p = parts[r]
                    assert L.orb_gpu_exchange_top2_device(hs[r], p[0].data_ptr(), p[1].data_ptr(), p[2].data_ptr(), nq,
                                                          out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(),
                                                          streams[r].cuda_stream) == 0
                    outs.append(out)
            else:
                launches = gpu.launch_count()
                for r, (lo, hi) in enumerate(((0, cut), (cut, ndb))):
                    out = torch.full((3, nq), -9, dtype=torch.int32, device="cuda")
                    assert L.orb_gpu_exchange_match_bruteforce_device(hs[r], dq.data_ptr(), nq, ddb[lo:hi].data_ptr(), hi - lo,
                                                                      lo, out[0].data_ptr(), out[1].data_ptr(),
                                                                      out[2].data_ptr(), streams[r].cuda_stream) == 0
                    outs.append(out)
                assert gpu.launch_count() - launches == 2      # one launch per rank: nothing but the match kernel
            torch.cuda.synchronize()
            for out in outs:
                assert np.array_equal(out.cpu().numpy(), want)
    finally:
        torch.cuda.synchronize()
        for r in range(2):
            L.orb_gpu_exchange_destroy(hs[r])
    assert gpu.launch_count() > before


def test_peer_memory_exchange_two_ranks_on_one_gpu(gpu, oracle):
    """orb_gpu_exchange_top2_device: both ranks of a world of 2 live in this process on one GPU (two streams, buffers
    wired with orb_gpu_exchange_connect_local); five calls in a row reuse the double-buffered gather slots."""
    import ctypes as C
    import torch
    L = gpu.lib()
    nq, ndb = 1000, 60000
    hs = (C.c_void_p * 2)()
    for r in range(2):
        h = C.c_void_p()
        assert L.orb_gpu_exchange_create(0, r, 2, 1024, C.byref(h)) == 0
        hs[r] = h
    assert L.orb_gpu_exchange_connect_local(hs, 2) == 0
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    try:
        for call in range(5):
            q, db = synth.make_match_set(nq, ndb, seed=40 + call)
            want = np.stack(oracle.match_bruteforce(q, db, threads=4))
            dq = torch.from_numpy(q).cuda()
            ddb = torch.from_numpy(db).cuda()
            cut = ndb // 3 + 17 * call
            parts, outs = [], []
            for r, (lo, hi) in enumerate(((0, cut), (cut, ndb))):
                part = torch.empty((3, nq), dtype=torch.int32, device="cuda")
                assert L.orb_gpu_match_bruteforce_device(dq.data_ptr(), nq, ddb[lo:hi].data_ptr(), hi - lo, lo,
                                                         part[0].data_ptr(), part[1].data_ptr(), part[2].data_ptr(),
                                                         None) == 0
                parts.append(part)
            torch.cuda.synchronize()
            for r in range(2):   # the two kernels wait for each other: they must be on different streams
                out = torch.empty((3, nq), dtype=torch.int32, device="cuda")
                p = parts[r]
                assert L.orb_gpu_exchange_top2_device(hs[r], p[0].data_ptr(), p[1].data_ptr(), p[2].data_ptr(), nq,
                                                      out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(),
                                                      streams[r].cuda_stream) == 0
                outs.append(out)
            torch.cuda.synchronize()
            for out in outs:
                assert np.array_equal(out.cpu().numpy(), want)
    finally:
        torch.cuda.synchronize()
        for r in range(2):
            L.orb_gpu_exchange_destroy(hs[r])
