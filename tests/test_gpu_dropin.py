"""The drop-in boundary, end to end in C++: the reference's own Frame constructors (src/Frame.cc compiled VERBATIM with
its own Frame.h) run on the product's drop-in ORBextractor.h (c_orb_slam_b200/csrc/adapter -> C ABI -> liborb_gpu.so),
exactly as a maintainer of the reference would build them (tests/cpp/Makefile).  Results must equal the oracle's."""
import ctypes as C
import os

import numpy as np
import pytest

from c_orb_slam_b200 import synth

LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp", "_build", "libdropin_frame.so")


def _lib():
    if not os.path.exists(LIB):
        pytest.skip("tests/cpp/_build/libdropin_frame.so not built (needs /root/reference)")
    return C.CDLL(LIB)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def test_dropin_library_exports():
    L = _lib()
    assert hasattr(L, "dropin_frame_mono") and hasattr(L, "dropin_frame_stereo")


@pytest.mark.gpu
@pytest.mark.parametrize("shape,nfeat", [((640, 480), 1000), ((752, 480), 1200)])
def test_reference_frame_constructor_on_the_dropin_extractor(gpu, oracle, shape, nfeat):
    L = _lib()
    W, H = shape
    img = synth.make_frame(W, H, 123)
    cap = nfeat + 24
    kps = np.zeros(cap, gpu.KP_DTYPE); desc = np.zeros((cap, 32), np.uint8)
    K4 = np.array([520.0, 516.0, 318.6, 255.3], np.float32)
    win = np.array([300.0, 200.0, 40.0, 0, 2], np.float32)
    area = np.zeros(cap, np.int32); n_area = C.c_int(); grid = np.zeros(4, np.float32)
    n = L.dropin_frame_mono(_p(img), W, H, nfeat, C.c_float(1.2), 8, 20, 7, _p(K4), _p(kps), _p(desc), cap, _p(win),
                            _p(area), C.byref(n_area), _p(grid))
    ko, do = oracle.Extractor(nfeat, 1.2, 8, 20, 7)(img)
    assert n == len(ko) and np.array_equal(kps[:n], ko) and np.array_equal(desc[:n], do)
    # the frame the reference built is searchable: its grid returns what the oracle's grid returns on those keypoints
    want = oracle.features_in_area(ko, tuple(grid), win[0], win[1], win[2], 0, 2)
    assert n_area.value == len(want) > 5 and np.array_equal(area[:n_area.value], want)
    assert grid[2] == np.float32(64) / np.float32(W) and grid[3] == np.float32(48) / np.float32(H)


@pytest.mark.gpu
def test_reference_stereo_frame_on_two_dropin_extractors(gpu, oracle):
    """Stereo constructor: two extractor instances called from two threads (src/Frame.cc:78-81), then the reference's
    own CPU ComputeStereoMatches on the pyramids the adapter put into mvImagePyramid."""
    L = _lib()
    W, H, nfeat = 1241, 376, 2000
    BF, FX = np.float32(386.1448), np.float32(718.856)
    left, right = synth.make_stereo_pair_subpixel(W, H, 31)
    cap = nfeat + 24
    kps = np.zeros(cap, gpu.KP_DTYPE); desc = np.zeros((cap, 32), np.uint8)
    ur = np.zeros(cap, np.float32); dp = np.zeros(cap, np.float32)
    K4 = np.array([FX, FX, 607.19, 185.2], np.float32)
    kr_g = np.zeros(cap, gpu.KP_DTYPE); dr_g = np.zeros((cap, 32), np.uint8); nr = C.c_int()
    sums = np.zeros(16, np.uint64)
    n = L.dropin_frame_stereo(_p(left), _p(right), W, H, nfeat, C.c_float(1.2), 8, 20, 7, _p(K4), C.c_float(BF), _p(kps),
                              _p(desc), cap, _p(ur), _p(dp), _p(kr_g), _p(dr_g), C.byref(nr), _p(sums))
    exl = oracle.Extractor(nfeat, 1.2, 8, 20, 7); exr = oracle.Extractor(nfeat, 1.2, 8, 20, 7)
    kl, dl = exl(left); kr, dr = exr(right)
    assert n == len(kl) and np.array_equal(kps[:n], kl) and np.array_equal(desc[:n], dl)
    assert nr.value == len(kr) and np.array_equal(kr_g[:nr.value], kr) and np.array_equal(dr_g[:nr.value], dr)

    def wsum(m):   # the position-weighted byte sum the harness takes of every mvImagePyramid level
        yy, xx = np.mgrid[0:m.shape[0], 0:m.shape[1]]
        return int((m.astype(np.uint64) * (1 + ((xx + 3 * yy) & 7)).astype(np.uint64)).sum())
    want = [wsum(exl.level_image(l)) for l in range(8)] + [wsum(exr.level_image(l)) for l in range(8)]
    assert [int(v) for v in sums] == want
    t = exl.tables()
    b = np.float32(BF / FX)
    max_d = np.float32(BF / b)
    bi, bd = oracle.stereo_match(kl, dl, kr, dr, H, t["scale"], 0.0, max_d)
    ur_o, dp_o, bad = oracle.stereo_refine(kl, kr, bi, bd, [exl.level_image(l) for l in range(8)],
                                           [exr.level_image(l) for l in range(8)], t["scale"], t["inv_scale"], 0.0, max_d, BF)
    assert bad == 0 and (ur_o >= 0).sum() > 0.2 * n
    assert np.array_equal(ur[:n].view(np.uint32), ur_o.view(np.uint32))
    assert np.array_equal(dp[:n].view(np.uint32), dp_o.view(np.uint32))


# ---- the matcher adapters (adapter/ORBmatcher_gpu.h) instantiated with the reference's real Frame ----------------
@pytest.mark.gpu
@pytest.mark.parametrize("th,stereo,seed", [(1.0, False, 1), (5.0, True, 3)])
def test_adapter_search_by_projection_equals_reference_on_real_frames(gpu, oracle, th, stereo, seed):
    """Reference ORBmatcher::SearchByProjection(Frame&, MapPoints, th) and the adapter's gpu::SearchByProjection on
    copies of one Frame built by the reference's constructor: same mvpMapPoints afterwards."""
    from test_gpu_matcher import _projection_case
    L = _lib()
    W, H = 640, 480
    kps, desc, proj, qdesc, u_right, blocked, grid, sf = _projection_case(oracle, seed, th, stereo)
    img = synth.make_frame(W, H, seed)          # the frame _projection_case extracted its keypoints from
    nk, nq = len(kps), len(proj)
    pg = np.zeros(nq, gpu.PROJ_DTYPE)
    for f in ("u", "v", "ur", "level", "view_cos", "valid", "has_obs"):
        pg[f] = proj[f]
    K4 = np.array([520.0, 516.0, 318.6, 255.3], np.float32)
    m_ref = np.zeros(nk + 8, np.int32); m_gpu = np.zeros(nk + 8, np.int32); n_ref = C.c_int(); n_gpu = C.c_int()
    N = L.dropin_search_by_projection(_p(img), W, H, 2000, _p(K4), _p(pg), _p(qdesc), nq,
                                      None if u_right is None else _p(u_right), _p(blocked), C.c_float(th),
                                      C.c_float(0.9), _p(m_ref), _p(m_gpu), C.byref(n_ref), C.byref(n_gpu))
    assert N == nk
    assert n_ref.value > 100 and n_ref.value == n_gpu.value and np.array_equal(m_ref[:N], m_gpu[:N])
    assert (m_ref[:N][blocked == 1] == -2).all()


@pytest.mark.gpu
@pytest.mark.parametrize("mode,seed", [("mono", 61), ("forward", 62), ("backward", 63)])
def test_adapter_last_frame_search_equals_reference_on_real_frames(gpu, oracle, mode, seed):
    """Reference SearchByProjection(CurrentFrame, LastFrame, th, bMono) and the adapter's SearchByProjectionLast, which
    projects with the reference's own expressions on the host and searches on the GPU."""
    from test_ref_matcher import CAM, _scene
    L = _lib()
    ex, rng, kps, desc, Tcw, src, Xw = _scene(oracle, seed)
    img = synth.make_frame(CAM["W"], CAM["H"], seed)
    nk, M = len(kps), len(src)
    dz = {"mono": 0.0, "forward": -0.5, "backward": 0.5}[mode]
    Tlw = Tcw.copy()
    Tlw[2, 3] -= np.float32(dz)
    stereo = mode != "mono"
    ur = np.where(rng.random(nk) < 0.6, kps["x"] - rng.uniform(1, 40, nk), -1).astype(np.float32) if stereo else None
    kl = np.zeros(M, gpu.KP_DTYPE)
    kl["octave"] = np.clip(kps["octave"][src] + rng.integers(-1, 2, M), 0, 7)
    rot = np.where(rng.random(M) < 0.85, rng.normal(40, 6, M), rng.uniform(0, 360, M))
    kl["angle"] = np.mod(kps["angle"][src] + rot, 360).astype(np.float32)
    has_mp = (rng.random(M) < 0.9).astype(np.uint8)
    outlier = (rng.random(M) < 0.05).astype(np.uint8)
    obs = (rng.random(M) < 0.9).astype(np.int32)
    mp_desc = synth.flip_bits(desc[src], rng.integers(0, 70, M), rng)
    holder = (rng.random(nk) < 0.05).astype(np.uint8)
    K4 = np.array([CAM["fx"], CAM["fy"], CAM["cx"], CAM["cy"]], np.float32)
    bf = np.float32(40.0)
    Tc = np.ascontiguousarray(Tcw, np.float32); Tl = np.ascontiguousarray(Tlw, np.float32)
    Xw = np.ascontiguousarray(Xw, np.float32)
    m_ref = np.zeros(nk + 8, np.int32); m_gpu = np.zeros(nk + 8, np.int32); n_ref = C.c_int(); n_gpu = C.c_int()
    N = L.dropin_search_last_frame(_p(img), CAM["W"], CAM["H"], 2000, _p(K4), C.c_float(bf), _p(Tc), _p(Tl),
                                   None if ur is None else _p(ur), _p(holder), _p(kl), M, _p(has_mp), _p(outlier), _p(Xw),
                                   _p(mp_desc), _p(obs), C.c_float(7.0), int(mode == "mono"), 1, _p(m_ref), _p(m_gpu),
                                   C.byref(n_ref), C.byref(n_gpu))
    assert N == nk
    assert n_ref.value > 100 and n_ref.value == n_gpu.value and np.array_equal(m_ref[:N], m_gpu[:N])


@pytest.mark.gpu
def test_adapter_stereo_equals_reference_cpu_stereo(gpu, oracle):
    """gpu::ComputeStereoMatches / gpu::StereoHammingMatches (adapter) against the reference's CPU
    Frame::ComputeStereoMatches, on the Frame the reference's stereo constructor built."""
    L = _lib()
    W, H, nfeat = 752, 480, 1200
    left, right = synth.make_stereo_pair_subpixel(W, H, 9)
    cap = nfeat + 24
    K4 = np.array([435.2, 435.2, 367.4, 252.2], np.float32)
    a = [np.zeros(cap, np.float32) for _ in range(4)]
    bi = np.zeros(cap, np.int32); bd = np.zeros(cap, np.int32)
    n = L.dropin_stereo_adapter(_p(left), _p(right), W, H, nfeat, _p(K4), C.c_float(47.9), _p(a[0]), _p(a[1]), _p(a[2]),
                                _p(a[3]), _p(bi), _p(bd))
    assert n > 1000 and (a[0][:n] >= 0).sum() > 0.2 * n
    assert np.array_equal(a[0][:n].view(np.uint32), a[2][:n].view(np.uint32))     # mvuRight
    assert np.array_equal(a[1][:n].view(np.uint32), a[3][:n].view(np.uint32))     # mvDepth
    matched = a[0][:n] >= 0
    assert (bi[:n][matched] >= 0).all() and (bd[:n][matched] < 75).all()           # thOrbDist = (TH_HIGH + TH_LOW) / 2
