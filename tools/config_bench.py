#!/usr/bin/env python
"""Secondary numbers for the five BASELINE.json configs on ONE GPU (the headline line is bench.py's).
Writes gpurun_out/configs.json; DESIGN.md / BASELINE.md quote it.  Product path only (C ABI); inputs synthetic."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import c_orb_slam_b200 as orb  # noqa: E402
from c_orb_slam_b200 import synth  # noqa: E402

L = orb.lib()
out = {}


def timed(fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


def device_batch(shape, nfeat, B, reps=10):
    """frames/s with the batch resident in HBM (device C ABI) and through the host-pointer call."""
    W, H = synth.SHAPES[shape]
    cap = nfeat + 24
    fr = synth.make_frames_fast(W, H, B, seed0=2000)
    h_fr = torch.from_numpy(fr).pin_memory()
    d_fr = h_fr.cuda()
    ex = orb.ORBextractor(nfeat, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
    d_k = torch.empty((B, cap, 28), dtype=torch.uint8, device="cuda")
    d_d = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(B, dtype=torch.int32, device="cuda")
    d_s = torch.zeros(1, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    dt = timed(lambda: ex.extract_batch_device(d_fr.data_ptr(), B, W, H, W, W * H, d_k.data_ptr(), d_d.data_ptr(),
                                               d_n.data_ptr(), d_s.data_ptr(), st), reps)
    assert int(d_s.item()) == 0
    hk = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
    hd = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
    hn = torch.zeros(B, dtype=torch.int32).pin_memory()

    def host():
        assert L.orb_gpu_extract_batch(ex._h, h_fr.data_ptr(), B, W, H, W, W * H, hk.data_ptr(), hd.data_ptr(), cap,
                                       hn.data_ptr()) == 0
    dth = timed(host, max(3, reps // 2))
    return {"shape": shape, "nfeatures": nfeat, "batch": B, "device_frames_per_s": B / dt, "device_ms": dt * 1e3,
            "host_api_frames_per_s": B / dth, "host_api_ms": dth * 1e3, "keypoints_per_frame": float(hn.sum()) / B}


# config 1: one TUM frame, 1000 features: latency of a single call
img = synth.make_frame(640, 480, 1000)
ex1 = orb.ORBextractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480, max_batch=1)
dt = timed(lambda: ex1(img), 200, warm=20)
out["config1_tum_single_frame"] = {"latency_us_host_api": dt * 1e6, "frames_per_s": 1 / dt, "keypoints": len(ex1(img)[0])}
out["config1_tum_batch512"] = device_batch("tum", 1000, 512)

# config 2: KITTI stereo pair, 2000 features each, extraction + stereo Hamming matching
W, H = synth.SHAPES["kitti"]
left, right = synth.make_stereo_pair(W, H, 31)
exl = orb.ORBextractor(2000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=1)
exr = orb.ORBextractor(2000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=1)
m = orb.ORBmatcher()
sf = np.asarray(exl.GetScaleFactors(), np.float32)


def stereo_pair():
    kl, dl = exl(left)
    kr, dr = exr(right)
    return m.stereo_match(kl, dl, kr, dr, H, sf, 0.0, np.float32(718.856))


dt = timed(stereo_pair, 50, warm=5)
bi, bd = stereo_pair()
out["config2_kitti_stereo_pair"] = {"ms_per_pair_host_api": dt * 1e3, "pairs_per_s": 1 / dt,
                                    "left_matches": int((bi >= 0).sum())}
BF, B_ = np.float32(386.1448), np.float32(386.1448) / np.float32(718.856)


def stereo_pair_full():
    """extraction of both images + the whole Frame::ComputeStereoMatches (Hamming + sub-pixel + median)"""
    kl, dl = exl(left)
    kr, dr = exr(right)
    return m.ComputeStereoMatches(exl, exr, kl, dl, kr, dr, BF, B_)


dt = timed(stereo_pair_full, 50, warm=5)
ur, dp = stereo_pair_full()
kl, dl = exl(left)
kr, dr = exr(right)
dt_sm = timed(lambda: m.ComputeStereoMatches(exl, exr, kl, dl, kr, dr, BF, B_), 50, warm=5)
out["config2_kitti_stereo_pair_full"] = {"ms_per_pair_host_api": dt * 1e3, "pairs_per_s": 1 / dt,
                                         "compute_stereo_matches_ms": dt_sm * 1e3, "depth_points": int((ur >= 0).sum())}

# batched stereo on the device: 128 pairs interleaved in one batch (left = 2i, right = 2i+1)
NP = 128
cap2 = 2000 + 24
lr = np.stack([left, right] * NP)
d_lr = torch.from_numpy(lr).cuda()
exb = orb.ORBextractor(2000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=2 * NP)
dk = torch.empty((2 * NP, cap2, 28), dtype=torch.uint8, device="cuda")
dd = torch.empty((2 * NP, cap2, 32), dtype=torch.uint8, device="cuda")
dn = torch.zeros(2 * NP, dtype=torch.int32, device="cuda")
ds = torch.zeros(1, dtype=torch.int32, device="cuda")
du = torch.empty((NP, cap2), dtype=torch.float32, device="cuda")
dz = torch.empty((NP, cap2), dtype=torch.float32, device="cuda")
dsad = torch.empty((NP, cap2), dtype=torch.int32, device="cuda")
st = torch.cuda.current_stream().cuda_stream


def stereo_batch(with_extract=True):
    if with_extract:
        exb.extract_batch_device(d_lr.data_ptr(), 2 * NP, W, H, W, W * H, dk.data_ptr(), dd.data_ptr(), dn.data_ptr(),
                                 ds.data_ptr(), st)
    assert L.orb_gpu_stereo_matches_device(exb._h, 0, 2, exb._h, 1, 2, NP, dk.data_ptr(), dd.data_ptr(), dn.data_ptr(),
                                           cap2, dk.data_ptr(), dd.data_ptr(), dn.data_ptr(), cap2, float(BF),
                                           float(B_), du.data_ptr(), dz.data_ptr(), dsad.data_ptr(), st) == 0


dt_all = timed(stereo_batch, 5, warm=2)
dt_sm = timed(lambda: stereo_batch(False), 10, warm=2)
out["config2_kitti_stereo_batch128_device"] = {"pairs_per_s": NP / dt_all, "ms_per_batch": dt_all * 1e3,
                                               "stereo_matches_ms_per_batch": dt_sm * 1e3,
                                               "depth_points_per_pair": float((du[0, :int(dn[0])] >= 0).sum())}
out["config2_kitti_batch256"] = device_batch("kitti", 2000, 256)

# config 3: EuRoC 1024 frames, 1200 features (one GPU's worth; frames shard across GPUs with no collective)
out["config3_euroc_batch1024"] = device_batch("euroc", 1200, 1024, reps=5)

# config 4: windowed SearchByProjection, 2000 map points x ~2000 keypoints, TH_HIGH, ratio 0.9
kps, desc = orb.ORBextractor(2000, 1.2, 8, 20, 7, max_width=640, max_height=480)(synth.make_frame(640, 480, 1))
rng = np.random.default_rng(1)
nk, M = len(kps), 2000
proj = np.zeros(M, orb.PROJ_DTYPE)
src = rng.integers(0, nk, M)
near = rng.random(M) < 0.75
proj["u"] = np.where(near, kps["x"][src] + rng.normal(0, 2, M), rng.uniform(0, 640, M)).astype(np.float32)
proj["v"] = np.where(near, kps["y"][src] + rng.normal(0, 2, M), rng.uniform(0, 480, M)).astype(np.float32)
proj["level"] = np.clip(kps["octave"][src] + rng.integers(0, 2, M), 0, 7)
proj["view_cos"] = rng.uniform(0.99, 1.0, M).astype(np.float32)
proj["valid"] = 1
proj["has_obs"] = 1
qdesc = synth.flip_bits(desc[src], rng.integers(0, 61, M), rng)
sf8 = np.asarray(orb.ORBextractor(2000, 1.2, 8, 20, 7, max_width=640, max_height=480).GetScaleFactors(), np.float32)
grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
mm = orb.ORBmatcher(0.9)


def sbp():
    frame = dict(kps=kps, desc=desc, u_right=None, grid=grid, scale_factors=sf8, blocked=np.zeros(nk, np.uint8))
    return mm.SearchByProjection(frame, dict(proj=proj, desc=qdesc), 1.0)


dt = timed(sbp, 50, warm=5)
out["config4_search_by_projection"] = {"ms_per_call_host_api": dt * 1e3, "map_points": M, "keypoints": int(nk),
                                       "matches": int(sbp()[0])}

# P1 / next rows: one call each on realistic sizes (host-pointer C ABI, allocations and copies included)
rng = np.random.default_rng(5)
tree = synth.make_vocabulary(10, 5, 1)
voc = orb.ORBVocabulary(tree)
dt = timed(lambda: voc.transform_features(desc, 4), 50, warm=5)
d_desc = torch.from_numpy(np.tile(desc, (64, 1))).cuda()
nf = d_desc.shape[0]
dw = torch.empty(nf, dtype=torch.int32, device="cuda"); dv = torch.empty(nf, dtype=torch.float64, device="cuda")
dnid = torch.empty(nf, dtype=torch.int32, device="cuda")
dt_dev = timed(lambda: L.orb_gpu_vocabulary_transform_device(voc._h, d_desc.data_ptr(), nf, 4, dw.data_ptr(),
                                                             dv.data_ptr(), dnid.data_ptr(), st), 20, warm=3)
out["n3_vocabulary_transform"] = {"nodes": int(len(tree["word_id"])), "features": int(nk),
                                  "ms_per_frame_host_api": dt * 1e3, "device_features_per_s": nf / dt_dev}
_, fv1 = voc.transform(desc, 2)
perm = rng.permutation(nk)
desc2 = synth.flip_bits(desc[perm], rng.integers(0, 45, nk), rng)
_, fv2 = voc.transform(desc2, 2)
good = np.ones(nk, bool)
mb = orb.ORBmatcher(0.7)
dt = timed(lambda: mb.SearchByBoW(dict(kps=kps, desc=desc, feat_vec=fv1), dict(kps=kps[perm], desc=desc2, feat_vec=fv2),
                                  good), 30, warm=3)
out["m6_search_by_bow"] = {"ms_per_call_host_api": dt * 1e3, "features": int(nk)}
prev = np.stack([kps["x"], kps["y"]], 1).astype(np.float32)
k2 = kps[perm].copy()
dt = timed(lambda: orb.ORBmatcher(0.9).SearchForInitialization(dict(kps=kps, desc=desc), dict(kps=k2, desc=desc2, grid=grid),
                                                              prev.copy(), 100), 30, warm=3)
out["m7_search_for_initialization"] = {"ms_per_call_host_api": dt * 1e3, "features": int(nk)}
lists = [synth.flip_bits(np.repeat(desc[i:i + 1], n, 0), rng.integers(0, 40, n), rng)
         for i, n in enumerate(rng.integers(2, 30, 2000))]
dt = timed(lambda: mb.ComputeDistinctiveDescriptors(lists), 10, warm=2)
out["n4_distinctive_descriptors"] = {"ms_per_2000_points_host_api": dt * 1e3}

os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", "configs.json"), "w") as f:
    json.dump(out, f, indent=1)
print(json.dumps(out, indent=1))
