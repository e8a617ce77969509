#!/usr/bin/env python
"""Generates tests/golden/*.npz: small known-answer fixtures that pin the oracle.

Run in the build container only (needs cv2 4.13.0 and /root/reference):
  * cv2_primitives.npz  -- cv2 outputs of resize / GaussianBlur / FAST / fastAtan2 on small inputs
  * ref_extract.npz     -- outputs of the reference's own ORBextractor.cc (compiled verbatim,
                           oracle/_ref/libref_orb_seq.so: list nodes from a bump arena, so the
                           (nKeys, pointer) sort equals the canonical creation-order tie rule)
                           and of the literal heap-order build (libref_orb.so) for the mismatch count
"""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from c_orb_slam_b200 import synth  # noqa: E402
from oracle import pyoracle as O  # noqa: E402

out = os.path.join(ROOT, "tests", "golden")
os.makedirs(out, exist_ok=True)
cv2.setNumThreads(1)
rng = np.random.default_rng(2024)

# ---- cv2 primitives --------------------------------------------------------------------------
img = synth.make_frame(160, 120, 9)
rnd = rng.integers(0, 256, (64, 80), dtype=np.uint8)
d = dict(img=img, rnd=rnd)
for name, (dw, dh) in dict(r133x100=(133, 100), r111x83=(111, 83), r40x30=(40, 30)).items():
    d["resize_img_" + name] = cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR)
d["resize_rnd_67x53"] = cv2.resize(rnd, (67, 53), interpolation=cv2.INTER_LINEAR)
d["blur_img"] = cv2.GaussianBlur(img, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
d["blur_rnd"] = cv2.GaussianBlur(rnd, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
d["border_rnd"] = cv2.copyMakeBorder(rnd, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
for th in (20, 7):
    fd = cv2.FastFeatureDetector_create(th, True)
    for nm, im in (("img", img), ("rnd", rnd), ("cell", np.ascontiguousarray(img[20:57, 30:66]))):
        kp = fd.detect(im)
        d["fast%d_%s" % (th, nm)] = np.array([(int(p.pt[0]), int(p.pt[1]), int(p.response)) for p in kp],
                                             np.int32).reshape(-1, 3)
ay = rng.integers(-(1 << 23), 1 << 23, 4096).astype(np.float32)
ax = rng.integers(-(1 << 23), 1 << 23, 4096).astype(np.float32)
ay[:8] = 0
ax[4:12] = 0
d["atan_y"], d["atan_x"] = ay, ax
d["atan_deg"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(ay, ax)], np.float32)
np.savez_compressed(os.path.join(out, "cv2_primitives.npz"), **d)

# ---- the reference extractor itself ------------------------------------------------------------
r = {}
for name, (W, H), nf, seed in (("tum", (640, 480), 1000, 1000), ("kitti", (1241, 376), 2000, 1001),
                               ("small", (320, 240), 500, 1002)):
    im = synth.make_frame(W, H, seed)
    ks, ds = O.RefExtractor(nf, 1.2, 8, 20, 7, seq=True)(im)
    kl, dl = O.RefExtractor(nf, 1.2, 8, 20, 7, seq=False)(im)
    r[name + "_args"] = np.array([W, H, nf, seed], np.int32)
    r[name + "_kps_seq"] = ks
    r[name + "_desc_seq"] = ds
    r[name + "_kps_heap"] = kl
    r[name + "_desc_heap"] = dl
rx = O.RefExtractor(1000, 1.2, 8, 20, 7)
t = rx.tables()
for k, v in t.items():
    r["tables_" + k] = v
np.savez_compressed(os.path.join(out, "ref_extract.npz"), **r)
for f in os.listdir(out):
    print(f, os.path.getsize(os.path.join(out, f)))
