"""GPU entry points against the reference's own src/ORBmatcher.cc compiled verbatim (oracle/_ref/libref_matcher.so,
which travels to the GPU box as a prebuilt file) -- no restated oracle in between.  Skipped where that library is absent."""
import numpy as np
import pytest

from test_gpu_bow_init import _two_views
from test_gpu_matcher import _projection_case, _to_gpu_queries
from test_ref_matcher import _cam_tuple, _scene
from c_orb_slam_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref(oracle):
    if not oracle.have_ref_matcher():
        pytest.skip("oracle/_ref/libref_matcher.so not built")
    return oracle


@pytest.mark.parametrize("th,stereo,seed", [(1.0, False, 21), (5.0, True, 22)])
def test_search_by_projection_gpu_equals_reference(gpu, ref, th, stereo, seed):
    kps, desc, proj, qdesc, u_right, blocked, grid, sf = _projection_case(ref, seed, th, stereo)
    n_r, m_r, b_r = ref.search_by_projection(proj, qdesc, kps, desc, u_right, grid, sf, th, 0.9, 100, blocked.copy(),
                                             impl="reference")
    frame = dict(kps=kps, desc=desc, u_right=u_right, grid=grid, scale_factors=sf, blocked=blocked.copy())
    pg = np.zeros(len(proj), gpu.PROJ_DTYPE)
    for f in ("u", "v", "ur", "level", "view_cos", "valid", "has_obs"):
        pg[f] = proj[f]
    n_g, m_g = gpu.ORBmatcher(0.9).SearchByProjection(frame, dict(proj=pg, desc=qdesc), th)
    assert n_r > 100 and n_g == n_r and np.array_equal(m_g, m_r) and np.array_equal(frame["blocked"], b_r)


def test_bow_and_initialization_gpu_equal_reference(gpu, ref):
    _, rng, (k1, d1, fv1), (k2, d2, fv2) = _two_views(ref, 31)
    good1 = rng.random(len(k1)) < 0.8
    good2 = rng.random(len(k2)) < 0.8
    n_r, m21_r = ref.search_by_bow_frame(fv1, fv2, k1, d1, good1, k2, d2, 0.7, True, impl="reference")
    n_g, _, m21_g = gpu.ORBmatcher(0.7, checkOri=True).SearchByBoW(
        dict(kps=k1, desc=d1, feat_vec=fv1), dict(kps=k2, desc=d2, feat_vec=fv2), good1)
    assert n_r > 200 and n_g == n_r and np.array_equal(m21_g, m21_r)
    n_r, m12_r = ref.search_by_bow_keyframes(fv1, fv2, k1, d1, good1, k2, d2, good2, 0.75, True, impl="reference")
    n_g, m12_g, _ = gpu.ORBmatcher(0.75, checkOri=True).SearchByBoW(
        dict(kps=k1, desc=d1, feat_vec=fv1), dict(kps=k2, desc=d2, feat_vec=fv2), good1, good2)
    assert n_r > 150 and n_g == n_r and np.array_equal(m12_g, m12_r)
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    prev_r, prev_g = prev.copy(), prev.copy()
    n_r, mi_r = ref.search_for_initialization(k1, d1, prev_r, k2, d2, grid, 100, 0.9, True, impl="reference")
    n_g, mi_g = gpu.ORBmatcher(0.9, checkOri=True).SearchForInitialization(
        dict(kps=k1, desc=d1), dict(kps=k2, desc=d2, grid=grid), prev_g, 100)
    assert n_r > 100 and n_g == n_r and np.array_equal(mi_g, mi_r) and np.array_equal(prev_g, prev_r)


@pytest.mark.parametrize("mode,seed", [("mono", 41), ("forward", 42)])
def test_last_frame_projection_gpu_equals_reference(gpu, ref, mode, seed):
    """The reference projects the last frame's points itself; the GPU gets the windows it searched."""
    ex, rng, kps, desc, Tcw, src, Xw = _scene(ref, seed)
    nk, M = len(kps), len(src)
    sf = ex.tables()["scale"].astype(np.float32)
    Tlw = Tcw.copy()
    Tlw[2, 3] -= np.float32(0.0 if mode == "mono" else -0.5)
    stereo = mode != "mono"
    ur_cur = np.where(rng.random(nk) < 0.6, kps["x"] - rng.uniform(1, 40, nk), -1).astype(np.float32) if stereo else None
    kps_last = np.zeros(M, ref.KP_DTYPE)
    kps_last["octave"] = np.clip(kps["octave"][src] + rng.integers(-1, 2, M), 0, 7)
    rot = np.where(rng.random(M) < 0.85, rng.normal(40, 6, M), rng.uniform(0, 360, M))
    kps_last["angle"] = np.mod(kps["angle"][src] + rot, 360).astype(np.float32)
    has_mp = rng.random(M) < 0.9
    outlier = rng.random(M) < 0.05
    mp_obs = (rng.random(M) < 0.9).astype(np.int32)
    mp_desc = synth.flip_bits(desc[src], rng.integers(0, 70, M), rng)
    blocked = (rng.random(nk) < 0.05).astype(np.uint8)
    cam = _cam_tuple(mbf=40.0, mb=0.1)
    n_r, m_r, win = ref.ref_search_last_frame(cam, Tcw, Tlw, kps, desc, ur_cur, sf, blocked, kps_last, has_mp, outlier,
                                              Xw, mp_desc, mp_obs, 7.0, mode == "mono", True)
    frame = dict(kps=kps, desc=desc, u_right=ur_cur, grid=(0.0, 0.0, cam[10], cam[11]), blocked=blocked.copy())
    n_g, m_g, _ = gpu.ORBmatcher(0.9, checkOri=True).SearchWindowed(frame, _to_gpu_queries(gpu, win), mp_desc, 100,
                                                                    1 if stereo else 0)
    assert n_r > 100 and n_g == n_r and np.array_equal(np.maximum(m_g, -1), m_r)


def test_compute_stereo_matches_gpu_equals_reference(gpu, oracle):
    """Whole Frame::ComputeStereoMatches on the device against the reference's own src/Frame.cc (verbatim build)."""
    if not oracle.have_ref_frame():
        pytest.skip("oracle/_ref/libref_frame.so not built")
    BF, FX = np.float32(386.1448), np.float32(718.856)
    W, H, nfeat = 1241, 376, 2000
    left, right = synth.make_stereo_pair_subpixel(W, H, 57)
    exl = gpu.ORBextractor(nfeat, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=1)
    exr = gpu.ORBextractor(nfeat, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=1)
    kl, dl = exl(left)
    kr, dr = exr(right)
    b = np.float32(BF / FX)
    ur, dp = gpu.ORBmatcher().ComputeStereoMatches(exl, exr, kl, dl, kr, dr, BF, b)
    oxl = oracle.Extractor(nfeat, 1.2, 8, 20, 7)
    oxr = oracle.Extractor(nfeat, 1.2, 8, 20, 7)
    oxl(left); oxr(right)
    t = oxl.tables()
    lv_l = [oxl.level_image(l) for l in range(8)]
    lv_r = [oxr.level_image(l) for l in range(8)]
    ur_r, dp_r, rc = oracle.ref_compute_stereo_matches(kl, dl, kr, dr, lv_l, lv_r, t["scale"], t["inv_scale"], b, BF)
    assert rc == 0 and (ur_r >= 0).sum() > 0.2 * len(kl)
    assert np.array_equal(ur.view(np.uint32), ur_r.view(np.uint32))
    assert np.array_equal(dp.view(np.uint32), dp_r.view(np.uint32))
