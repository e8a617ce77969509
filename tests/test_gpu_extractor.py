"""GPU parity of the extractor against the CPU oracle, through the C ABI (pytest -m gpu).

Bar: pyramid pixels, FAST candidates (order included), keypoint set/order/octave/response,
descriptors: bit-exact.  Angles: bit-exact in practice (same float ops), asserted <= 1e-3 deg.
"""
import numpy as np
import pytest

from c_orb_slam_b200 import synth

pytestmark = pytest.mark.gpu

CONFIGS = [("tum", 1000), ("euroc", 1200), ("kitti", 2000)]


def _compare_frame(ex_o, ex_g, img, frame=0, kps_g=None, desc_g=None):
    ko, do = ex_o(img)
    if kps_g is None:
        kps_g, desc_g = ex_g(img)
    for l in range(8):
        assert np.array_equal(ex_o.level_image(l), ex_g.pyramid_level(l, frame)), "pyramid level %d" % l
        co, cg = ex_o.candidates(l), ex_g.candidates(l, frame)
        assert co.shape == cg.shape and np.array_equal(co, cg), "FAST candidates level %d" % l
        bo = ex_o.level_image(l, blurred=True)
        if bo is not None:
            assert np.array_equal(bo, ex_g.pyramid_level(l, frame, blurred=True)), "blur level %d" % l
    assert len(ko) == len(kps_g)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(ko[f], kps_g[f]), f
    dang = np.abs(ko["angle"].astype(np.float64) - kps_g["angle"].astype(np.float64))
    assert dang.max() <= 1e-3, "angle tolerance 1e-3 deg (north_star)"
    nbad = int((do != desc_g).any(axis=1).sum())
    return dict(n=len(ko), angle_bitdiff=int((ko["angle"] != kps_g["angle"]).sum()), desc_rows_diff=nbad)


@pytest.mark.parametrize("shape,nfeat", CONFIGS)
def test_single_frame_parity(gpu, oracle, shape, nfeat):
    W, H = synth.SHAPES[shape]
    ex_o = oracle.Extractor(nfeat, 1.2, 8, 20, 7)
    ex_g = gpu.ORBextractor(nfeat, 1.2, 8, 20, 7, max_width=W, max_height=H)
    t = ex_o.tables()
    assert np.array_equal(t["scale"], ex_g.GetScaleFactors())
    assert np.array_equal(t["inv_scale"], ex_g.GetInverseScaleFactors())
    assert np.array_equal(t["sigma2"], ex_g.GetScaleSigmaSquares())
    assert np.array_equal(t["inv_sigma2"], ex_g.GetInverseScaleSigmaSquares())
    q, u = ex_g.features_per_level()
    assert np.array_equal(q, t["quota"]) and np.array_equal(u, t["umax"])
    assert ex_g.GetLevels() == 8 and abs(ex_g.GetScaleFactor() - 1.2) < 1e-6
    for seed in (1000, 1001):
        r = _compare_frame(ex_o, ex_g, synth.make_frame(W, H, seed))
        assert r["desc_rows_diff"] == 0 and r["angle_bitdiff"] == 0, r


def test_batch_matches_single(gpu, oracle):
    W, H = synth.SHAPES["euroc"]
    B = 6
    frames = synth.make_frames(W, H, B, seed0=2000)
    ex_g = gpu.ORBextractor(1200, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=4)  # forces 2 chunks
    n, kps, desc = ex_g.extract_batch(frames)
    ex_o = oracle.Extractor(1200, 1.2, 8, 20, 7)
    for f in range(B):
        ko, do = ex_o(frames[f])
        assert n[f] == len(ko)
        assert np.array_equal(kps[f, :n[f]], ko) and np.array_equal(desc[f, :n[f]], do)


def test_strided_input_and_reuse(gpu, oracle):
    W, H = 640, 480
    big = np.zeros((H, W + 37), np.uint8)
    big[:, :W] = synth.make_frame(W, H, 77)
    ex_g = gpu.ORBextractor(1000, 1.2, 8, 20, 7, max_width=800, max_height=600)
    ex_o = oracle.Extractor(1000, 1.2, 8, 20, 7)
    k1, d1 = ex_g(big[:, :W])
    ko, do = ex_o(np.ascontiguousarray(big[:, :W]))
    assert np.array_equal(k1, ko) and np.array_equal(d1, do)
    # same handle, different geometry, then back
    small = synth.make_frame(400, 300, 5)
    k2, d2 = ex_g(small)
    ko2, do2 = ex_o(small)
    assert np.array_equal(k2, ko2) and np.array_equal(d2, do2)
    k3, d3 = ex_g(big[:, :W])
    assert np.array_equal(k3, ko) and np.array_equal(d3, do)


def test_edge_cases(gpu, oracle):
    ex_g = gpu.ORBextractor(500, 1.2, 8, 20, 7, max_width=640, max_height=480)
    k, d = ex_g(np.empty((0, 0), np.uint8))            # empty image: silent return (:1046)
    assert len(k) == 0 and d.shape == (0, 32)
    flat = np.full((480, 640), 128, np.uint8)           # no corners anywhere
    k, d = ex_g(flat)
    assert len(k) == 0
    ko, _ = oracle.Extractor(500, 1.2, 8, 20, 7)(flat)
    assert len(ko) == 0
    with pytest.raises(gpu.OrbGpuError):                # level 7 would have no FAST cell
        ex_g(np.zeros((100, 100), np.uint8))
    with pytest.raises(gpu.OrbGpuError):                # larger than max_width
        ex_g(np.zeros((480, 700), np.uint8))
    # low-texture image: exercises the minThFAST retry and under-full quotas
    rng = np.random.default_rng(3)
    soft = (128 + 12 * rng.standard_normal((480, 640))).clip(0, 255).astype(np.uint8)
    k, d = ex_g(soft)
    ko, do = oracle.Extractor(500, 1.2, 8, 20, 7)(soft)
    assert np.array_equal(k, ko) and np.array_equal(d, do)


def test_non_default_parameters(gpu, oracle):
    img = synth.make_frame(752, 480, 4242)
    for args in ((300, 1.5, 4, 30, 10), (3000, 1.1, 10, 15, 5)):
        ex_g = gpu.ORBextractor(*args, max_width=752, max_height=480)
        ex_o = oracle.Extractor(*args)
        k, d = ex_g(img)
        ko, do = ex_o(img)
        assert np.array_equal(k, ko) and np.array_equal(d, do), args


def test_device_api_aliased_and_unaligned(gpu, oracle):
    """orb_gpu_extract_batch_device: level 0 aliases aligned caller frames, repacks unaligned ones."""
    import torch
    for (W, H), nf in (((640, 480), 1000), ((1241, 376), 2000)):   # 1241-px rows are not 4-byte aligned
        B = 3
        frames = synth.make_frames(W, H, B, seed0=500)
        d_frames = torch.from_numpy(frames).cuda()
        cap = nf + 24
        d_kps = torch.zeros((B, cap, 28), dtype=torch.uint8, device="cuda")
        d_desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
        d_n = torch.zeros(B, dtype=torch.int32, device="cuda")
        d_st = torch.ones(1, dtype=torch.int32, device="cuda")
        ex = gpu.ORBextractor(nf, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            ex.extract_batch_device(d_frames.data_ptr(), B, W, H, W, W * H, d_kps.data_ptr(), d_desc.data_ptr(),
                                    d_n.data_ptr(), d_st.data_ptr(), s.cuda_stream)
        s.synchronize()
        assert int(d_st.item()) == 0
        n = d_n.cpu().numpy()
        kps = d_kps.cpu().numpy().view(gpu.KP_DTYPE).reshape(B, cap)
        desc = d_desc.cpu().numpy()
        ex_o = oracle.Extractor(nf, 1.2, 8, 20, 7)
        for f in range(B):
            ko, do = ex_o(frames[f])
            assert n[f] == len(ko) and np.array_equal(kps[f, :n[f]], ko) and np.array_equal(desc[f, :n[f]], do)
        assert np.array_equal(ex.pyramid_level(0, 1), frames[1])


@pytest.mark.parametrize("W,H,scale,nlevels", [(641, 480, 1.2, 8), (642, 479, 1.2, 8), (643, 481, 1.3, 6),
                                               (1241, 376, 1.2, 8), (800, 600, 2.0, 4), (640, 480, 1.05, 8)])
def test_pyramid_levels_pixel_exact_for_odd_widths_and_scales(gpu, oracle, W, H, scale, nlevels):
    """Every pyramid level, pixel for pixel, for widths of every residue mod 4 (unaligned rows are repacked, the
    four-column resize stores whole words into the row padding) and for scale factors up to the limit of 2."""
    img = synth.make_frame(W, H, 77)
    ex_g = gpu.ORBextractor(400, scale, nlevels, 20, 7, max_width=W, max_height=H)
    ex_o = oracle.Extractor(400, scale, nlevels, 20, 7)
    k, d = ex_g(img)
    ko, do = ex_o(img)
    for l in range(nlevels):
        assert np.array_equal(ex_g.pyramid_level(l), ex_o.level_image(l)), l
    assert np.array_equal(k, ko) and np.array_equal(d, do)


@pytest.mark.parametrize("W,H,scale,nlevels,B", [(640, 480, 1.2, 8, 3), (225, 150, 1.1, 3, 2), (449, 131, 1.2, 4, 5),
                                                 (1241, 376, 1.2, 8, 2), (752, 480, 1.2, 8, 1), (400, 300, 1.08, 12, 2),
                                                 (1920, 1080, 1.5, 5, 1)])
def test_blurred_levels_equal_gaussian_blur_of_the_levels(gpu, oracle, W, H, scale, nlevels, B):
    """Every blurred level of every frame of a batch == GaussianBlur 7x7 sigma 2 REFLECT_101 (src/ORBextractor.cc:1085-1086)
    of the unblurred level, pixel for pixel: the tensor-core kernel's stacked frames (REFLECT_101 rows folded into the
    Toeplitz matrix at every frame boundary), column chunks with one or two image edges, chunks of a few columns, levels
    shorter than a strip, more than eight levels (two launches) and 1241-px rows (repacked level 0)."""
    frames = synth.make_frames(W, H, B, seed0=91)
    ex = gpu.ORBextractor(300, scale, nlevels, 20, 7, max_width=W, max_height=H, max_batch=B)
    ex.extract_batch(frames)
    for f in range(B):
        for l in range(nlevels):
            want = oracle.gaussian_blur7(ex.pyramid_level(l, f))
            assert np.array_equal(ex.pyramid_level(l, f, blurred=True), want), (f, l)


def test_blur_of_frames_tma_cannot_address(gpu, oracle):
    """Device frames whose rows are 4-byte but not 16-byte aligned and frames with a gap between them: level 0 goes
    through the byte-SIMD blur kernel, the other levels through the tensor-core kernel; all equal the oracle."""
    import torch
    W, H, B = 636, 200, 3
    stride, fstride = 644, 644 * 200 + 644
    buf = torch.zeros(B * fstride + 64, dtype=torch.uint8, device="cuda")
    frames = synth.make_frames(W, H, B, seed0=17)
    host = np.zeros(B * fstride + 64, np.uint8)
    for f in range(B):
        host[f * fstride:f * fstride + H * stride].reshape(H, stride)[:, :W] = frames[f]
    buf.copy_(torch.from_numpy(host))
    ex = gpu.ORBextractor(300, 1.2, 6, 20, 7, max_width=W, max_height=H, max_batch=B)
    cap = ex.capacity
    d_kps = torch.empty((B, cap, 28), dtype=torch.uint8, device="cuda")
    d_desc = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.empty(B, dtype=torch.int32, device="cuda")
    ex.extract_batch_device(buf.data_ptr(), B, W, H, stride, fstride, d_kps.data_ptr(), d_desc.data_ptr(), d_n.data_ptr())
    torch.cuda.synchronize()
    n = d_n.cpu().numpy()
    kps = d_kps.cpu().numpy().view(gpu.KP_DTYPE).reshape(B, cap)
    desc = d_desc.cpu().numpy()
    ex_o = oracle.Extractor(300, 1.2, 6, 20, 7)
    for f in range(B):
        for l in range(6):
            assert np.array_equal(ex.pyramid_level(l, f, blurred=True), oracle.gaussian_blur7(ex.pyramid_level(l, f))), (f, l)
        ko, do = ex_o(frames[f])
        assert n[f] == len(ko) and np.array_equal(kps[f, :n[f]], ko) and np.array_equal(desc[f, :n[f]], do)


@pytest.mark.parametrize("W,H,scale,nlevels,B", [(640, 480, 1.2, 8, 9), (752, 480, 1.2, 8, 8), (1241, 376, 1.2, 8, 8),
                                                 (643, 231, 1.3, 5, 11), (400, 300, 1.08, 10, 8), (1920, 1080, 1.5, 4, 8),
                                                 (800, 600, 1.85, 3, 8)])
@pytest.mark.parametrize("variant", [0, 1])
def test_batch_pyramid_levels_pixel_exact(gpu, oracle, W, H, scale, nlevels, B, variant):
    """Every level of every frame of a batch == cv::resize of the level below (src/ORBextractor.cc:1107-1132), pixel
    for pixel, for both resize forms: the byte-SIMD kernel (default) and the tensor-core form (stacked frames, strips that
    cross frame boundaries, partial column chunks, weight matrices rebuilt per chunk, scale factors up to its limit of ~1.85)."""
    frames = synth.make_frames(W, H, B, seed0=123)
    ex = gpu.ORBextractor(300, scale, nlevels, 20, 7, max_width=W, max_height=H, max_batch=B)
    assert gpu.lib().orb_gpu_set_resize_variant(variant) == 0
    try:
        ex.extract_batch(frames)
    finally:
        gpu.lib().orb_gpu_set_resize_variant(0)
    for f in range(B):
        prev = frames[f]
        for l in range(1, nlevels):
            got = ex.pyramid_level(l, f)
            want = oracle.resize_linear(prev, got.shape[1], got.shape[0])
            if not np.array_equal(got, want):
                ys, xs = np.nonzero(got != want)
                raise AssertionError("frame %d level %d: %d mismatches, rows %d..%d cols %d..%d, first (y=%d,x=%d) got %d want %d" % (
                    f, l, len(ys), ys.min(), ys.max(), xs.min(), xs.max(), ys[0], xs[0], got[ys[0], xs[0]], want[ys[0], xs[0]]))
            prev = got
