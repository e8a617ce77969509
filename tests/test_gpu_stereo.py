"""N1 parity: Frame::ComputeStereoMatches (src/Frame.cc:466-640) on the device -- Hamming stage, 11x11 SAD
sub-pixel refinement on the extractors' resident pyramids, median filter -- against the oracle.  mvuRight and
mvDepth are float results of a fixed operation sequence: compared bit for bit."""
import ctypes as C

import numpy as np
import pytest

from c_orb_slam_b200 import synth

pytestmark = pytest.mark.gpu

BF, FX = np.float32(386.1448), np.float32(718.856)   # KITTI 00-02 (Examples/Stereo/KITTI00-02.yaml)


def _oracle_stereo(oracle, left, right, nfeat):
    exl = oracle.Extractor(nfeat, 1.2, 8, 20, 7)
    exr = oracle.Extractor(nfeat, 1.2, 8, 20, 7)
    kl, dl = exl(left)
    kr, dr = exr(right)
    t = exl.tables()
    b = np.float32(BF / FX)
    max_d = np.float32(BF / b)
    bi, bd = oracle.stereo_match(kl, dl, kr, dr, left.shape[0], t["scale"], 0.0, max_d)
    lv_l = [exl.level_image(l) for l in range(8)]
    lv_r = [exr.level_image(l) for l in range(8)]
    ur, dp, bad = oracle.stereo_refine(kl, kr, bi, bd, lv_l, lv_r, t["scale"], t["inv_scale"], 0.0, max_d, BF)
    return (kl, dl, kr, dr), ur, dp, bad, b


@pytest.mark.parametrize("shape,nfeat,seed,subpixel", [
    ((640, 480), 1000, 3, True), ((1241, 376), 2000, 31, True), ((752, 480), 1200, 5, False)])
def test_compute_stereo_matches(gpu, oracle, shape, nfeat, seed, subpixel):
    W, H = shape
    mk = synth.make_stereo_pair_subpixel if subpixel else synth.make_stereo_pair
    left, right = mk(W, H, seed)
    (kl_o, dl_o, kr_o, dr_o), ur_o, dp_o, bad, b = _oracle_stereo(oracle, left, right, nfeat)
    exl = gpu.ORBextractor(nfeat, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=1)
    exr = gpu.ORBextractor(nfeat, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=1)
    kl, dl = exl(left)
    kr, dr = exr(right)
    assert np.array_equal(kl, kl_o) and np.array_equal(kr, kr_o) and np.array_equal(dl, dl_o)
    ur, dp = gpu.ORBmatcher().ComputeStereoMatches(exl, exr, kl, dl, kr, dr, BF, b)
    assert bad == 0
    n_ok = int((ur_o >= 0).sum())
    assert n_ok > 0.2 * len(kl)
    assert np.array_equal(ur.view(np.uint32), ur_o.view(np.uint32))
    assert np.array_equal(dp.view(np.uint32), dp_o.view(np.uint32))
    if subpixel:   # the parabola really produces fractional coordinates
        frac = np.abs(ur[ur >= 0] / 1.0 - np.rint(ur[ur >= 0]))
        assert (frac > 0.05).mean() > 0.3
    # properties of :611-621 and :623-639
    ok = ur >= 0
    assert ((kl["x"][ok] - ur[ok]) > 0).all() and ((kl["x"][ok] - ur[ok]) < BF / b).all()
    assert np.allclose(dp[ok], BF / (kl["x"][ok] - ur[ok]), rtol=1e-6)


def test_stereo_matches_device_batch(gpu, oracle):
    """Four stereo pairs interleaved in one batch (frames 2i = left, 2i+1 = right) through the device entry
    point; every pair must equal the oracle."""
    import torch
    W, H, nfeat, npairs = 640, 480, 800, 4
    cap = nfeat + 24
    pairs = [synth.make_stereo_pair_subpixel(W, H, 40 + i) for i in range(npairs)]
    frames = np.stack([im for p in pairs for im in p])
    ex = gpu.ORBextractor(nfeat, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=2 * npairs)
    d_fr = torch.from_numpy(frames).cuda()
    d_k = torch.zeros((2 * npairs, cap, 28), dtype=torch.uint8, device="cuda")
    d_d = torch.zeros((2 * npairs, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(2 * npairs, dtype=torch.int32, device="cuda")
    d_s = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    ex.extract_batch_device(d_fr.data_ptr(), 2 * npairs, W, H, W, W * H, d_k.data_ptr(), d_d.data_ptr(),
                            d_n.data_ptr(), d_s.data_ptr(), st)
    d_u = torch.empty((npairs, cap), dtype=torch.float32, device="cuda")
    d_z = torch.empty((npairs, cap), dtype=torch.float32, device="cuda")
    d_sad = torch.empty((npairs, cap), dtype=torch.int32, device="cuda")
    b = np.float32(BF / FX)
    rc = gpu.lib().orb_gpu_stereo_matches_device(ex._h, 0, 2, ex._h, 1, 2, npairs, d_k.data_ptr(), d_d.data_ptr(),
                                                 d_n.data_ptr(), cap, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(),
                                                 cap, float(BF), float(b), d_u.data_ptr(), d_z.data_ptr(),
                                                 d_sad.data_ptr(), st)
    assert rc == 0
    torch.cuda.synchronize()
    assert int(d_s.item()) == 0
    n = d_n.cpu().numpy()
    for i, (left, right) in enumerate(pairs):
        _, ur_o, dp_o, bad, _ = _oracle_stereo(oracle, left, right, nfeat)
        nl = int(n[2 * i])
        assert nl == len(ur_o) and bad == 0
        assert np.array_equal(d_u[i, :nl].cpu().numpy().view(np.uint32), ur_o.view(np.uint32))
        assert np.array_equal(d_z[i, :nl].cpu().numpy().view(np.uint32), dp_o.view(np.uint32))


def test_stereo_argument_checks(gpu):
    ex = gpu.ORBextractor(500, 1.2, 8, 20, 7, max_width=320, max_height=240, max_batch=1)
    other = gpu.ORBextractor(500, 1.2, 6, 20, 7, max_width=320, max_height=240, max_batch=1)
    k = np.zeros(1, gpu.KP_DTYPE); d = np.zeros((1, 32), np.uint8)
    m = gpu.ORBmatcher()
    with pytest.raises(gpu.OrbGpuError):      # nothing extracted yet
        m.ComputeStereoMatches(ex, ex, k, d, k, d, BF, 0.5)
    img = synth.make_frame(320, 240, 1)
    ex(img); other(img)
    with pytest.raises(gpu.OrbGpuError):      # different pyramids
        m.ComputeStereoMatches(ex, other, k, d, k, d, BF, 0.5)
    ur, dp = m.ComputeStereoMatches(ex, ex, k[:0], d[:0], k, d, BF, 0.5)
    assert len(ur) == 0 and len(dp) == 0
