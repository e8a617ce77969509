"""M6 / M7 parity on the GPU: SearchByBoW (KF,Frame) and (KF,KF), SearchForTriangulation and
SearchForInitialization through the C ABI against the oracle's restatement of src/ORBmatcher.cc:159-288,
:405-520, :522-655, :657-823.  Integer results: bit-exact."""
import numpy as np
import pytest

from c_orb_slam_b200 import synth

pytestmark = pytest.mark.gpu


def _two_views(oracle, seed, n_nodes=97, dup=250):
    """Frame 2 = the features of frame 1 in another order, moved a little, with noisy descriptors; both get a
    DBoW2-style feature vector (node -> feature list) in which a feature and its twin mostly share the node.
    Some frame-1 descriptors are near copies of each other, so several sources want one target."""
    ex = oracle.Extractor(1500, 1.2, 8, 20, 7)
    kps1, desc1 = ex(synth.make_frame(640, 480, 50 + seed))
    rng = np.random.default_rng(seed)
    n = len(kps1)
    a = rng.integers(0, n, dup)
    b = (a + rng.integers(1, 4, dup)) % n
    desc1 = desc1.copy()
    desc1[b] = synth.flip_bits(desc1[a], rng.integers(0, 6, dup), rng)
    perm = rng.permutation(n)
    kps2 = kps1[perm].copy()
    kps2["x"] = (kps2["x"] + rng.normal(0, 2.0, n)).astype(np.float32)
    kps2["y"] = (kps2["y"] + rng.normal(0, 1.2, n)).astype(np.float32)
    kps2["angle"] = np.mod(kps2["angle"] + np.where(rng.random(n) < 0.85, rng.normal(25, 5, n), rng.uniform(0, 360, n)),
                           360).astype(np.float32)
    desc2 = synth.flip_bits(desc1[perm], rng.integers(0, 75, n), rng)
    node1 = rng.integers(0, n_nodes, n)
    node1[b] = node1[a]
    node2 = np.where(rng.random(n) < 0.9, node1[perm], rng.integers(0, n_nodes + 5, n))
    fv1, fv2 = {}, {}
    for i in range(n):
        fv1.setdefault(int(node1[i]), []).append(i)
        fv2.setdefault(int(node2[i]), []).append(i)
    return ex, rng, (kps1, desc1, fv1), (kps2, desc2, fv2)


@pytest.mark.parametrize("check_ori,ratio,seed", [(True, 0.7, 1), (False, 0.9, 2), (True, 0.75, 3)])
def test_search_by_bow_keyframe_frame(gpu, oracle, check_ori, ratio, seed):
    _, rng, (k1, d1, fv1), (k2, d2, fv2) = _two_views(oracle, seed)
    good1 = rng.random(len(k1)) < 0.8
    n_o, m21_o = oracle.search_by_bow_frame(fv1, fv2, k1, d1, good1, k2, d2, ratio, check_ori)
    n_g, m12_g, m21_g = gpu.ORBmatcher(ratio, checkOri=check_ori).SearchByBoW(
        dict(kps=k1, desc=d1, feat_vec=fv1), dict(kps=k2, desc=d2, feat_vec=fv2), good1)
    assert n_o > 200
    assert n_g == n_o and np.array_equal(m21_g, m21_o)
    hit = np.flatnonzero(m21_g >= 0)
    assert np.array_equal(m12_g[m21_g[hit]], hit) and (m12_g >= 0).sum() == len(hit)


@pytest.mark.parametrize("check_ori,ratio,seed", [(True, 0.75, 4), (False, 0.8, 5)])
def test_search_by_bow_keyframes(gpu, oracle, check_ori, ratio, seed):
    _, rng, (k1, d1, fv1), (k2, d2, fv2) = _two_views(oracle, seed)
    good1 = rng.random(len(k1)) < 0.8
    good2 = rng.random(len(k2)) < 0.8
    # distances of exactly TH_LOW must be rejected here (strict '<', :598) and accepted in the other form
    n_o, m12_o = oracle.search_by_bow_keyframes(fv1, fv2, k1, d1, good1, k2, d2, good2, ratio, check_ori)
    n_g, m12_g, m21_g = gpu.ORBmatcher(ratio, checkOri=check_ori).SearchByBoW(
        dict(kps=k1, desc=d1, feat_vec=fv1), dict(kps=k2, desc=d2, feat_vec=fv2), good1, good2)
    assert n_o > 150
    assert n_g == n_o and np.array_equal(m12_g, m12_o)
    assert good2[m12_g[m12_g >= 0]].all() and good1[np.flatnonzero(m12_g >= 0)].all()


def test_bow_threshold_is_strict_only_between_keyframes(gpu, oracle):
    """One source, one candidate at distance exactly TH_LOW = 50: matched by (KF, Frame), not by (KF, KF)."""
    rng = np.random.default_rng(0)
    d1 = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    d2 = synth.flip_bits(d1, [50], rng)
    k = np.zeros(1, gpu.KP_DTYPE)
    fv = {7: [0]}
    m = gpu.ORBmatcher(0.9, checkOri=False)
    f = dict(kps=k, desc=d1, feat_vec=fv), dict(kps=k, desc=d2, feat_vec=fv)
    assert m.SearchByBoW(f[0], f[1], [1])[0] == 1
    assert m.SearchByBoW(f[0], f[1], [1], [1])[0] == 0
    assert oracle.search_by_bow_frame(fv, fv, k, d1, [1], k, d2, 0.9, False)[0] == 1
    assert oracle.search_by_bow_keyframes(fv, fv, k, d1, [1], k, d2, [1], 0.9, False)[0] == 0


@pytest.mark.parametrize("stereo,only_stereo,check_ori,seed", [
    (False, False, True, 6), (True, False, True, 7), (True, True, False, 8)])
def test_search_for_triangulation(gpu, oracle, stereo, only_stereo, check_ori, seed):
    ex, rng, (k1, d1, fv1), (k2, d2, fv2) = _two_views(oracle, seed)
    n1, n2 = len(k1), len(k2)
    has1 = rng.random(n1) < 0.3
    has2 = rng.random(n2) < 0.3
    ur1 = ur2 = None
    if stereo:
        ur1 = np.where(rng.random(n1) < 0.5, k1["x"] - 10, -1).astype(np.float32)
        ur2 = np.where(rng.random(n2) < 0.5, k2["x"] - 10, -1).astype(np.float32)
    # sideways motion: epipolar lines nearly horizontal, plus a small perturbation so that every term counts
    F12 = (np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32) +
           rng.normal(0, 2e-6, (3, 3)).astype(np.float32))
    sf = ex.tables()["scale"].astype(np.float32)
    sigma2 = (sf * sf).astype(np.float32)
    epi = (np.float32(k2["x"][5]), np.float32(k2["y"][5]))   # an epipole inside the image: its gate bites
    n_o, m12_o = oracle.search_for_triangulation(fv1, fv2, k1, d1, ur1, has1, k2, d2, ur2, has2, F12, epi[0], epi[1],
                                                 sf, sigma2, only_stereo, check_ori)
    n_g, m12_g = gpu.ORBmatcher(0.6, checkOri=check_ori).SearchForTriangulation(
        dict(kps=k1, desc=d1, feat_vec=fv1, u_right=ur1),
        dict(kps=k2, desc=d2, feat_vec=fv2, u_right=ur2, scale_factors=sf, level_sigma2=sigma2),
        F12, epi, has1, has2, only_stereo)
    assert n_o > (40 if only_stereo else 150)
    assert n_g == n_o and np.array_equal(m12_g, m12_o)
    # the epipolar gate must bite: with a huge sigma2 more pairs survive
    n_loose, _ = oracle.search_for_triangulation(fv1, fv2, k1, d1, ur1, has1, k2, d2, ur2, has2, F12, epi[0], epi[1],
                                                 sf, sigma2 * 1e6, only_stereo, False)
    n_tight, _ = oracle.search_for_triangulation(fv1, fv2, k1, d1, ur1, has1, k2, d2, ur2, has2, F12, epi[0], epi[1],
                                                 sf, sigma2, only_stereo, False)
    assert n_loose > n_tight


@pytest.mark.parametrize("check_ori,window,seed", [(True, 100, 9), (False, 100, 10), (True, 30, 11)])
def test_search_for_initialization(gpu, oracle, check_ori, window, seed):
    _, rng, (k1, d1, _), (k2, d2, _) = _two_views(oracle, seed, dup=400)
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)   # first call: vbPrevMatched = F1 positions
    n_o, m12_o = oracle.search_for_initialization(k1, d1, prev, k2, d2, grid, window, 0.9, check_ori)
    prev_g = prev.copy()
    n_g, m12_g = gpu.ORBmatcher(0.9, checkOri=check_ori).SearchForInitialization(
        dict(kps=k1, desc=d1), dict(kps=k2, desc=d2, grid=grid), prev_g, window)
    assert n_o > 100
    assert n_g == n_o and np.array_equal(m12_g, m12_o)
    assert (k1["octave"][m12_g >= 0] == 0).all() and (k2["octave"][m12_g[m12_g >= 0]] == 0).all()
    hit = m12_g >= 0
    assert len(np.unique(m12_g[hit])) == hit.sum()            # one source per target after the steal-back
    assert np.array_equal(prev_g[hit, 0], k2["x"][m12_g[hit]]) and np.array_equal(prev_g[~hit], prev[~hit])
    # second call from the updated positions (Tracking::MonocularInitialization calls it once per frame)
    n_o2, m12_o2 = oracle.search_for_initialization(k1, d1, prev_g, k2, d2, grid, window, 0.9, check_ori)
    n_g2, m12_g2 = gpu.ORBmatcher(0.9, checkOri=check_ori).SearchForInitialization(
        dict(kps=k1, desc=d1), dict(kps=k2, desc=d2, grid=grid), prev_g.copy(), window)
    assert n_g2 == n_o2 and np.array_equal(m12_g2, m12_o2)


def test_bucket_search_edge_cases(gpu):
    m = gpu.ORBmatcher(0.9)
    k = np.zeros(3, gpu.KP_DTYPE)
    d = np.zeros((3, 32), np.uint8)
    empty = dict(kps=k[:0], desc=d[:0], feat_vec={})
    full = dict(kps=k, desc=d, feat_vec={1: [0, 1, 2]})
    assert m.SearchByBoW(empty, full, [])[0] == 0
    n, m12, m21 = m.SearchByBoW(full, empty, [1, 1, 1])
    assert n == 0 and (m12 == -1).all() and len(m21) == 0
    n, m12, _ = m.SearchByBoW(full, dict(kps=k, desc=d, feat_vec={2: [0, 1, 2]}), [1, 1, 1])   # no common node
    assert n == 0 and (m12 == -1).all()
    q = np.array([(5, 0, 1)], gpu.BUCKET_DTYPE)   # source index out of range
    cand = np.zeros(1, np.int32)
    import ctypes as C
    out = np.empty(3, np.int32)
    nm = C.c_int()
    rc = gpu.lib().orb_gpu_search_by_bow(q.ctypes.data, 1, cand.ctypes.data, 1, k.ctypes.data, d.ctypes.data, 3,
                                         k.ctypes.data, d.ctypes.data, None, 3, 0.9, 50, 0, 0, out.ctypes.data,
                                         out.ctypes.data, C.byref(nm))
    assert rc == -1


def test_hand_cases_steal_back_and_ties(gpu, oracle):
    """The hand cases of tests/test_oracle.py through the C ABI: steal-back only when strictly closer, the last
    equal distance wins in SearchForTriangulation, a claimed frame feature is invisible in SearchByBoW."""
    rng = np.random.default_rng(3)
    target = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    grid = (0.0, 0.0, np.float32(0.1), np.float32(0.1))
    k2 = np.zeros(1, gpu.KP_DTYPE); k2["x"], k2["y"] = 100.0, 100.0
    k1 = np.zeros(3, gpu.KP_DTYPE); k1["x"], k1["y"] = [100.0, 101.0, 99.0], [100.0, 100.0, 101.0]
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    m = gpu.ORBmatcher(0.9, checkOri=False)
    for flips, want in (([10, 5, 20], [-1, 0, -1]), ([10, 10, 12], [0, -1, -1]), ([30, 20, 10], [-1, -1, 0]),
                        ([51, 50, 50], [-1, 0, -1])):
        d1 = synth.flip_bits(np.repeat(target, 3, 0), flips, rng)
        n, m12 = m.SearchForInitialization(dict(kps=k1, desc=d1), dict(kps=k2, desc=target, grid=grid), prev.copy())
        assert m12.tolist() == want and n == 1
    base = target
    d2 = synth.flip_bits(np.repeat(base, 3, 0), [7, 7, 9], rng)
    q1 = np.zeros(1, gpu.KP_DTYPE); q1["x"], q1["y"] = 50.0, 60.0
    t2 = np.zeros(3, gpu.KP_DTYPE); t2["x"], t2["y"] = [10.0, 20.0, 30.0], 60.0
    F = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
    one = np.ones(8, np.float32)
    kf1 = dict(kps=q1, desc=base, feat_vec={1: [0]}, u_right=None)
    kf2 = dict(kps=t2, desc=d2, feat_vec={1: [0, 1, 2]}, u_right=None, scale_factors=one, level_sigma2=one)
    n, m12 = m.SearchForTriangulation(kf1, kf2, F, (1e6, 1e6), [0], [0, 0, 0])
    assert n == 1 and m12.tolist() == [1]
    t2["y"][1] = 63.0
    n, m12 = m.SearchForTriangulation(kf1, kf2, F, (1e6, 1e6), [0], [0, 0, 0])
    assert m12.tolist() == [0]
    dd2 = synth.flip_bits(np.repeat(base, 2, 0), [0, 30], rng)
    dd1 = synth.flip_bits(np.repeat(base, 2, 0), [2, 1], rng)
    kk = np.zeros(2, gpu.KP_DTYPE)
    fv = {3: [0, 1]}
    n, m12, m21 = m.SearchByBoW(dict(kps=kk, desc=dd1, feat_vec=fv), dict(kps=kk, desc=dd2, feat_vec=fv), [1, 1])
    assert n == 2 and m21.tolist() == [0, 1]
    n, m12, m21 = gpu.ORBmatcher(0.05, checkOri=False).SearchByBoW(
        dict(kps=kk, desc=dd1, feat_vec=fv), dict(kps=kk, desc=dd2, feat_vec=fv), [1, 0])
    assert n == 0


def test_distinctive_descriptors(gpu, oracle):
    """N4: MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-307) batched over map points."""
    rng = np.random.default_rng(12)
    lists = []
    for n in [0, 1, 2, 3, 4, 5, 8, 33, 64, 100, 257, 1024] + rng.integers(2, 40, 300).tolist():
        base = rng.integers(0, 256, (1, 32), dtype=np.uint8)
        d = synth.flip_bits(np.repeat(base, n, 0), rng.integers(0, 60, n), rng) if n else np.zeros((0, 32), np.uint8)
        if n > 4:
            d[rng.integers(0, n)] = d[rng.integers(0, n)]   # equal rows: ties between medians
        lists.append(d)
    want = oracle.distinctive_descriptors(lists)
    got = gpu.ORBmatcher().ComputeDistinctiveDescriptors(lists)
    assert np.array_equal(got, want)
    assert got[0] == -1 and got[1] == 0 and (got[2:] >= 0).all()
    # independent restatement for one point
    d = lists[9].astype(np.uint8)
    dist = np.unpackbits(d[:, None, :] ^ d[None, :, :], axis=2).sum(2)
    med = np.sort(dist, 1)[:, int(0.5 * (len(d) - 1))]
    assert got[9] == int(np.argmin(med))
    with pytest.raises(gpu.OrbGpuError):
        gpu.ORBmatcher().ComputeDistinctiveDescriptors([np.zeros((1025, 32), np.uint8)])
