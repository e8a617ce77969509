"""The restated oracle against the reference's own src/Frame.cc compiled VERBATIM with its own include/Frame.h
(oracle/_ref/libref_frame.so, oracle/Makefile target `refframe`): Frame::ComputeStereoMatches (M10 + N1) and the
feature grid (M11).  CPU only; skipped where the reference build is absent."""
import numpy as np
import pytest

from c_orb_slam_b200 import synth

BF, FX = np.float32(386.1448), np.float32(718.856)   # Examples/Stereo/KITTI00-02.yaml


@pytest.fixture(scope="module")
def ref(oracle):
    if not oracle.have_ref_frame():
        pytest.skip("oracle/_ref/libref_frame.so not built (needs /root/reference)")
    oracle.ref_frame()
    return oracle


@pytest.mark.parametrize("shape,nfeat,seed,subpixel", [
    ((640, 480), 1000, 3, True), ((1241, 376), 2000, 31, True), ((752, 480), 1200, 5, False), ((640, 480), 600, 9, True)])
def test_compute_stereo_matches_equals_reference(ref, shape, nfeat, seed, subpixel):
    W, H = shape
    left, right = (synth.make_stereo_pair_subpixel if subpixel else synth.make_stereo_pair)(W, H, seed)
    exl = ref.Extractor(nfeat, 1.2, 8, 20, 7)
    exr = ref.Extractor(nfeat, 1.2, 8, 20, 7)
    kl, dl = exl(left)
    kr, dr = exr(right)
    t = exl.tables()
    b = np.float32(BF / FX)
    max_d = np.float32(BF / b)
    lv_l = [exl.level_image(l) for l in range(8)]
    lv_r = [exr.level_image(l) for l in range(8)]
    bi, bd = ref.stereo_match(kl, dl, kr, dr, H, t["scale"], 0.0, max_d)
    ur_o, dp_o, bad = ref.stereo_refine(kl, kr, bi, bd, lv_l, lv_r, t["scale"], t["inv_scale"], 0.0, max_d, BF)
    ur_r, dp_r, rc = ref.ref_compute_stereo_matches(kl, dl, kr, dr, lv_l, lv_r, t["scale"], t["inv_scale"], b, BF)
    assert rc == 0 and bad == 0
    assert (ur_r >= 0).sum() > 0.2 * len(kl)
    assert np.array_equal(ur_o.view(np.uint32), ur_r.view(np.uint32))
    assert np.array_equal(dp_o.view(np.uint32), dp_r.view(np.uint32))


def test_features_in_area_equals_reference(ref):
    ex = ref.Extractor(2000, 1.2, 8, 20, 7)
    kps, _ = ex(synth.make_frame(640, 480, 77))
    rng = np.random.default_rng(3)
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    total = 0
    for k in range(400):
        x, y = np.float32(rng.uniform(-30, 670)), np.float32(rng.uniform(-30, 510))
        r = np.float32(rng.choice([2.5, 7.0, 15.0, 40.0, 100.0]))
        lo, hi = [(-1, -1), (0, 0), (2, -1), (0, 3), (1, 2), (3, 5)][k % 6]
        a = ref.features_in_area(kps, grid, x, y, r, lo, hi)
        b = ref.ref_features_in_area(kps, grid, x, y, r, lo, hi)
        assert np.array_equal(a, b)           # same candidates in the same order
        total += len(b)
    assert total > 5000
    # keypoints on the grid border and outside the image (undistorted coordinates may leave it, :387-389)
    edge = kps[:64].copy()
    edge["x"] = np.linspace(-12, 652, 64).astype(np.float32)
    edge["y"] = np.float32(479.9)
    for x in (0.0, 320.0, 639.9):
        a = ref.features_in_area(edge, grid, np.float32(x), np.float32(478.0), np.float32(30.0), -1, -1)
        b = ref.ref_features_in_area(edge, grid, np.float32(x), np.float32(478.0), np.float32(30.0), -1, -1)
        assert np.array_equal(a, b)
