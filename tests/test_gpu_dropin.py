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
