#!/usr/bin/env python
"""A few matcher calls on realistic sizes (for ncu launch lists): SearchByProjection, SearchWindowed,
SearchByBoW, SearchForInitialization, ComputeStereoMatches, vocabulary transform."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import c_orb_slam_b200 as orb  # noqa: E402
from c_orb_slam_b200 import synth  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
ex = orb.ORBextractor(2000, 1.2, 8, 20, 7, max_width=640, max_height=480)
kps, desc = ex(synth.make_frame(640, 480, 1))
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
sf8 = np.asarray(ex.GetScaleFactors(), np.float32)
grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
mm = orb.ORBmatcher(0.9)
for _ in range(reps):
    frame = dict(kps=kps, desc=desc, u_right=None, grid=grid, scale_factors=sf8, blocked=np.zeros(nk, np.uint8))
    mm.SearchByProjection(frame, dict(proj=proj, desc=qdesc), 1.0)
res = orb.GpuFrame().upload(kps, desc, None, grid)
for _ in range(reps):   # the same search against a resident frame
    mm.SearchByProjection(dict(resident=res, scale_factors=sf8, blocked=np.zeros(nk, np.uint8)), dict(proj=proj, desc=qdesc), 1.0)
perm = rng.permutation(nk)
desc2 = synth.flip_bits(desc[perm], rng.integers(0, 45, nk), rng)
k2 = kps[perm].copy()
prev = np.stack([kps["x"], kps["y"]], 1).astype(np.float32)
for _ in range(reps):
    mm.SearchForInitialization(dict(kps=kps, desc=desc), dict(kps=k2, desc=desc2, grid=grid), prev.copy(), 100)
tree = synth.make_vocabulary(10, 4, 1)
voc = orb.ORBVocabulary(tree)
_, fv1 = voc.transform(desc, 2)
_, fv2 = voc.transform(desc2, 2)
for _ in range(reps):
    mm.SearchByBoW(dict(kps=kps, desc=desc, feat_vec=fv1), dict(kps=k2, desc=desc2, feat_vec=fv2), np.ones(nk, bool))
W, H = synth.SHAPES["kitti"]
left, right = synth.make_stereo_pair_subpixel(W, H, 31)
exl = orb.ORBextractor(2000, 1.2, 8, 20, 7, max_width=W, max_height=H)
exr = orb.ORBextractor(2000, 1.2, 8, 20, 7, max_width=W, max_height=H)
kl, dl = exl(left)
kr, dr = exr(right)
for _ in range(reps):
    mm.ComputeStereoMatches(exl, exr, kl, dl, kr, dr, 386.1448, 386.1448 / 718.856)
print("done")
