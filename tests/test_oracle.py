"""Pins the CPU oracle (not gpu): against committed cv2 4.13.0 / reference fixtures
(tests/golden, made by tools/gen_golden.py), against cv2 live when importable, and against the
reference's own ORBextractor.cc (oracle/_ref) when it is built."""
import os

import numpy as np
import pytest

from c_orb_slam_b200 import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def prim():
    return np.load(os.path.join(GOLD, "cv2_primitives.npz"))


@pytest.fixture(scope="module")
def refgold():
    return np.load(os.path.join(GOLD, "ref_extract.npz"))


def test_generator_is_stable(prim):
    assert np.array_equal(synth.make_frame(160, 120, 9), prim["img"])


def test_resize_blur_border_fixtures(oracle, prim):
    img, rnd = prim["img"], prim["rnd"]
    for name, (dw, dh) in dict(r133x100=(133, 100), r111x83=(111, 83), r40x30=(40, 30)).items():
        assert np.array_equal(oracle.resize_linear(img, dw, dh), prim["resize_img_" + name])
    assert np.array_equal(oracle.resize_linear(rnd, 67, 53), prim["resize_rnd_67x53"])
    assert np.array_equal(oracle.gaussian_blur7(img), prim["blur_img"])
    assert np.array_equal(oracle.gaussian_blur7(rnd), prim["blur_rnd"])
    assert np.array_equal(oracle.copy_make_border(rnd, 19, 19, 19, 19), prim["border_rnd"])


def test_fast_and_atan_fixtures(oracle, prim):
    img = prim["img"]
    for th in (20, 7):
        for nm, im in (("img", img), ("rnd", prim["rnd"]), ("cell", np.ascontiguousarray(img[20:57, 30:66]))):
            assert np.array_equal(oracle.fast9_nms(im, th), prim["fast%d_%s" % (th, nm)]), (th, nm)
    assert np.array_equal(oracle.fast_atan2(prim["atan_y"], prim["atan_x"]), prim["atan_deg"])


def test_constructor_tables(oracle, refgold):
    t = oracle.Extractor(1000, 1.2, 8, 20, 7).tables()
    for k in ("scale", "inv_scale", "sigma2", "inv_sigma2"):
        assert np.array_equal(t[k], refgold["tables_" + k])
    assert t["quota"].tolist() == [217, 181, 151, 126, 105, 87, 73, 60]
    assert t["umax"].tolist() == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert oracle.Extractor(1200, 1.2, 8, 20, 7).tables()["quota"].tolist() == [261, 217, 181, 151, 126, 105, 87, 72]
    assert oracle.Extractor(2000, 1.2, 8, 20, 7).tables()["quota"].tolist() == [434, 362, 302, 251, 209, 175, 145, 122]


@pytest.mark.parametrize("name", ["tum", "kitti", "small"])
def test_extractor_equals_reference_fixture(oracle, refgold, name):
    """Oracle == reference ORBextractor.cc (verbatim build, creation-order list-node addresses):
    keypoints, order, angles and descriptors bit for bit.  Also counts how many keypoints the
    literal heap-address tie-break changes (reported)."""
    W, H, nf, seed = refgold[name + "_args"].tolist()
    k, d = oracle.Extractor(nf, 1.2, 8, 20, 7)(synth.make_frame(W, H, seed))
    ks, ds = refgold[name + "_kps_seq"], refgold[name + "_desc_seq"]
    assert len(k) == len(ks)
    assert k.tobytes() == ks.tobytes() and np.array_equal(d, ds)
    kh = refgold[name + "_kps_heap"]
    a = {(p["x"], p["y"], p["octave"]) for p in k}
    b = {(p["x"], p["y"], p["octave"]) for p in kh}
    frac = len(a - b) / max(1, len(a))
    print("%s: heap-order tie-break changes %d of %d keypoints" % (name, len(a - b), len(a)))
    assert frac < 0.05


def test_extractor_equals_reference_live(oracle):
    if not oracle.have_ref():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    for (W, H), nf, seed in (((752, 480), 1200, 7), ((400, 300), 800, 8)):
        img = synth.make_frame(W, H, seed)
        ex, rx = oracle.Extractor(nf, 1.2, 8, 20, 7), oracle.RefExtractor(nf, 1.2, 8, 20, 7, seq=True)
        k, d = ex(img)
        kr, dr = rx(img)
        assert k.tobytes() == kr.tobytes() and np.array_equal(d, dr)
        for l in range(8):
            assert np.array_equal(ex.level_image(l), rx.level_image(l))


def test_primitives_vs_cv2_live(oracle):
    cv2 = pytest.importorskip("cv2")
    if not cv2.__version__.startswith("4.13"):
        pytest.skip("oracle arithmetic is pinned to cv2 4.13.0")
    cv2.setNumThreads(1)
    img = synth.make_frame(752, 480, 3)
    prev = img
    for (w, h) in ((627, 400), (522, 333), (435, 278)):
        ref = cv2.resize(prev, (w, h), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(oracle.resize_linear(prev, w, h), ref)
        assert np.array_equal(oracle.gaussian_blur7(prev), cv2.GaussianBlur(prev, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101))
        prev = ref
    fd = cv2.FastFeatureDetector_create(20, True)
    kp = fd.detect(img)
    r = np.array([(int(p.pt[0]), int(p.pt[1]), int(p.response)) for p in kp], np.int32).reshape(-1, 3)
    assert np.array_equal(oracle.fast9_nms(img, 20), r)


def test_quadtree_properties(oracle):
    rng = np.random.default_rng(5)
    for n, N in ((0, 50), (1, 50), (40, 50), (3000, 217), (3000, 1)):
        xy = rng.permutation(608 * 448)[:n]
        xyr = np.stack([xy % 608, xy // 608, rng.integers(7, 200, n)], 1).astype(np.float32)
        sel = oracle.distribute_quadtree(xyr, 16, 624, 16, 464, N)
        assert len(set(sel.tolist())) == len(sel)
        if n >= max(N, 4):
            assert N <= len(sel) <= N + 3
        else:
            assert len(sel) <= n and (n == 0 or len(sel) >= 1)  # a no-growth pass ends the loop early (:669-672)


def test_matcher_oracle_small(oracle):
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, 32, dtype=np.uint8)
    b = rng.integers(0, 256, 32, dtype=np.uint8)
    ref = int(np.unpackbits(a ^ b).sum())
    assert oracle.descriptor_distance(a, b) == ref
    q, db = synth.make_match_set(50, 700, seed=3)
    idx, best, second = oracle.match_bruteforce(q, db, threads=2)
    dist = np.unpackbits(q[:, None, :] ^ db[None, :, :], axis=2).sum(2)
    assert np.array_equal(best, dist.min(1)) and np.array_equal(idx, dist.argmin(1))
    assert np.array_equal(second, np.sort(dist, 1)[:, 1])


def _kp(oracle, xs, ys, octave=0, angle=0.0):
    k = np.zeros(len(xs), oracle.KP_DTYPE)
    k["x"], k["y"], k["octave"], k["angle"] = xs, ys, octave, angle
    return k


def test_oracle_initialization_steal_back(oracle):
    """Hand cases of SearchForInitialization (src/ORBmatcher.cc:405-520): a later source takes a target only when
    strictly closer (:444-445); the earlier source is un-matched (:463-467); level > 0 sources never match."""
    rng = np.random.default_rng(3)
    target = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    grid = (0.0, 0.0, np.float32(0.1), np.float32(0.1))
    k2 = _kp(oracle, [100.0], [100.0])
    k1 = _kp(oracle, [100.0, 101.0, 99.0], [100.0, 100.0, 101.0])
    prev = np.stack([k1["x"], k1["y"]], 1)
    for flips, want in (([10, 5, 20], [-1, 0, -1]), ([10, 10, 12], [0, -1, -1]), ([30, 20, 10], [-1, -1, 0]),
                        ([51, 50, 50], [-1, 0, -1])):
        d1 = synth.flip_bits(np.repeat(target, 3, 0), flips, rng)
        n, m12 = oracle.search_for_initialization(k1, d1, prev, k2, target, grid, 100, 0.9, False)
        assert m12.tolist() == want and n == 1
    k1["octave"] = [1, 0, 2]
    d1 = synth.flip_bits(np.repeat(target, 3, 0), [0, 40, 0], rng)
    n, m12 = oracle.search_for_initialization(k1, d1, prev, k2, target, grid, 100, 0.9, False)
    assert m12.tolist() == [-1, 0, -1]


def test_oracle_bow_claims_and_ratio(oracle):
    """SearchByBoW hand cases (:159-288, :522-655): a claimed frame feature is invisible to later sources of the
    node (:209); the ratio test is strict and uses 256 when the node holds one candidate (:230)."""
    rng = np.random.default_rng(4)
    base = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    d2 = synth.flip_bits(np.repeat(base, 2, 0), [0, 30], rng)          # two frame features in node 3
    d1 = synth.flip_bits(np.repeat(base, 2, 0), [2, 1], rng)           # both sources prefer frame feature 0
    k = _kp(oracle, [0.0, 0.0], [0.0, 0.0])
    fv = {3: [0, 1]}
    n, m21 = oracle.search_by_bow_frame(fv, fv, k, d1, [1, 1], k, d2, 0.9, False)
    # source 0 takes feature 0 (2 < 0.9 * ~30); source 1 then only sees feature 1, alone in the node: vs 256
    assert n == 2 and m21.tolist() == [0, 1]
    n, m21 = oracle.search_by_bow_frame(fv, fv, k, d1, [0, 1], k, d2, 0.9, False)
    assert n == 1 and m21.tolist() == [1, -1]
    n, m21 = oracle.search_by_bow_frame(fv, {4: [0, 1]}, k, d1, [1, 1], k, d2, 0.9, False)
    assert n == 0
    # ratio: best 2 vs second ~30 passes at 0.9 but not at 0.05
    n, _ = oracle.search_by_bow_frame(fv, fv, k, d1, [1, 0], k, d2, 0.05, False)
    assert n == 0


def test_oracle_triangulation_tie_goes_to_last(oracle):
    """SearchForTriangulation :737 skips only dist > bestDist, so an equal distance later in the node wins."""
    rng = np.random.default_rng(5)
    base = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    d2 = synth.flip_bits(np.repeat(base, 3, 0), [7, 7, 9], rng)
    k1 = _kp(oracle, [50.0], [60.0])
    k2 = _kp(oracle, [10.0, 20.0, 30.0], [60.0, 60.0, 60.0])
    F = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
    one = np.ones(8, np.float32)
    n, m12 = oracle.search_for_triangulation({1: [0]}, {1: [0, 1, 2]}, k1, base, None, [0], k2, d2, None, [0, 0, 0], F,
                                             1e6, 1e6, one, one, False, False)
    assert n == 1 and m12.tolist() == [1]
    k2["y"][1] = 63.0   # off the epipolar line: 9 > 3.84
    n, m12 = oracle.search_for_triangulation({1: [0]}, {1: [0, 1, 2]}, k1, base, None, [0], k2, d2, None, [0, 0, 0], F,
                                             1e6, 1e6, one, one, False, False)
    assert m12.tolist() == [0]
