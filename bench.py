#!/usr/bin/env python
"""Benchmark of the ORB front-end hot path (BASELINE.json metric: "ORB extract frames/s (TUM 640x480,
1000 kp); Hamming Gpairs/s").

  python bench.py --gpus N --steps K --warmup W            # this repo (CUDA, sm_100a)
  python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU code on host cores

One step = ORBextractor::operator() over one batch of synthetic TUM-shaped frames (TUM1.yaml
settings: 1000 features, 8 levels, 1.2, iniThFAST 20, minThFAST 7).  `value` is frames/s with the
frames resident in HBM (device C-ABI, CUDA events on the launching stream); `e2e` is the same work
through the host-pointer C-ABI call with pinned host buffers, H2D/D2H inside the timed region.
Frames shard across ranks (independent units, no collective): weak scaling.  Secondary objects in
the same line: `hamming` (brute-force 4096 x 10M top-2, database sharded across ranks, exact merge
over NVLink peer memory; with its own cpu_baseline), `parity` (outputs of the timed runs checked
against the reference's own code: keypoints / descriptors of bench frames, the count of keypoints
the literal heap-order tie-break changes, an oracle slice of the Hamming run), `latency_us_batch1`
(configs[0] as written: one frame per call) and `configs` (KITTI stereo pair, EuRoC 1024 frames
split over the ranks, windowed SearchByProjection, SearchByBoW -- each with the reference's CPU
code timed beside it at N = 1).  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SETTINGS = dict(nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7)  # Examples/Monocular/TUM1.yaml:30-43
W, H = 640, 480
ALGO_BYTES_PER_FRAME = 4481474      # pyramid r/w + FAST read + blur r/w + K*60
P_PIXELS = 950532                   # sum of level pixels, TUM
METRIC = "ORB extract frames/s (TUM 640x480, 1000 kp); Hamming Gpairs/s @1-8 GPU"


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


CLOCK_SAMPLE_S = float(os.environ.get("ORB_BENCH_CLOCK_MS", "2")) * 1e-3


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU while the timed region runs (NVML)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                     nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                     nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                     nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
            while not self.stop_flag:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(CLOCK_SAMPLE_S)
        except Exception as e:  # NVML missing: report nothing rather than guess
            self.reasons.add("nvml_unavailable:%s" % type(e).__name__)

    def result(self):
        self.stop_flag = True
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def cpu_reference_extract(frames, threads):
    """Times the reference's CPU extractor on `frames` with `threads` host threads.
    Returns (seconds, kind, total_keypoints)."""
    from oracle import pyoracle as O
    O.lib()
    s = SETTINGS
    if os.path.exists(O.REF_PATH):
        t0 = time.perf_counter()
        counts = O.ref_extract_batch(frames, s["nfeatures"], s["scale"], s["nlevels"], s["ini_th"], s["min_th"], threads)
        return time.perf_counter() - t0, "reference", int(counts.sum())
    t0 = time.perf_counter()
    counts, _, _ = O.extract_batch(frames, s["nfeatures"], s["scale"], s["nlevels"], s["ini_th"], s["min_th"],
                                   threads, want_output=False)
    return time.perf_counter() - t0, "port", int(counts.sum())


def cv2_primitive_bound(frames, threads):
    """Sanity bound on the speed of the restated OpenCV primitives: the pixel stages of one extraction (7 resizes,
    FAST over every level at iniThFAST, 8 Gaussian blurs) through the real OpenCV of this image (cv2: SIMD / IPP,
    `threads` threads).  No cells, retry, quadtree, orientation or descriptors, so a frame here costs LESS than a
    frame of ORBextractor::operator() over real OpenCV; frames/s, or None when cv2 is missing."""
    try:
        import cv2
    except Exception:
        return None
    s = SETTINGS
    cv2.setNumThreads(threads)
    fast = cv2.FastFeatureDetector_create(s["ini_th"], True)
    t0 = time.perf_counter()
    for img in frames:
        sf = 1.0
        level = img
        for l in range(s["nlevels"]):
            if l:
                sf *= s["scale"]
                level = cv2.resize(level, (int(round(W / sf)), int(round(H / sf))), interpolation=cv2.INTER_LINEAR)
            fast.detect(level)
            cv2.GaussianBlur(level, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
    return len(frames) / (time.perf_counter() - t0)


# ---------------------------------------------------------------------------------------------------------------
# parity of the timed runs (checked against the reference's own code, reported in the JSON line)
# ---------------------------------------------------------------------------------------------------------------
def parity_extraction(frames_np, kps, desc, n):
    """kps [F, cap] (KP_DTYPE), desc [F, cap, 32], n [F]: what the timed device run left for the first F bench
    frames.  Checked against the reference's ORBextractor.cc compiled verbatim: libref_orb_seq.so (list nodes from a
    bump arena: equal-sized quadtree nodes are taken newest first, the canonical tie rule the GPU implements) --
    every field, angle bits and descriptor rows -- and against libref_orb.so as built (ties of src/ORBextractor.cc:684
    follow heap addresses): the keypoints that differ there are counted per level, not hidden."""
    from oracle import pyoracle as O
    s = SETTINGS
    F = len(frames_np)
    args = (s["nfeatures"], s["scale"], s["nlevels"], s["ini_th"], s["min_th"])
    have_ref = O.have_ref()
    canon = O.RefExtractor(*args, seq=True) if have_ref else O.Extractor(*args)
    literal = O.RefExtractor(*args, seq=False) if have_ref else None
    total = kp_bad = angle_bad = desc_bad = count_bad = 0
    lit_diff = [0] * s["nlevels"]
    lit_total = 0
    geom = ("x", "y", "size", "response", "octave", "class_id")
    for f in range(F):
        k, d = kps[f, :n[f]], desc[f, :n[f]]
        kr, dr = canon(frames_np[f])
        total += len(kr)
        if len(k) != len(kr):
            count_bad += 1
            kp_bad += abs(len(k) - len(kr))
            continue
        same = np.ones(len(k), bool)
        for fld in geom:
            same &= k[fld] == kr[fld]
        kp_bad += int((~same).sum())
        angle_bad += int((k["angle"].view(np.uint32) != kr["angle"].view(np.uint32))[same].sum())
        desc_bad += int((d != dr).any(axis=1)[same].sum())
        if literal is not None:
            kl, _ = literal(frames_np[f])
            lit_total += len(kl)
            have = {(float(p["x"]), float(p["y"]), int(p["octave"])) for p in kl}
            for p in k:
                if (float(p["x"]), float(p["y"]), int(p["octave"])) not in have:
                    lit_diff[int(p["octave"])] += 1
    out = {"frames_checked": F, "keypoints_checked": total,
           "against": "reference src/ORBextractor.cc compiled verbatim, creation-order tie rule (oracle/_ref/libref_orb_seq.so)"
                      if have_ref else "restated oracle (oracle/_ref not built on this box)",
           "frames_with_count_mismatch": count_bad, "keypoint_mismatches": kp_bad,
           "angle_bit_mismatches": angle_bad, "descriptor_rows_differing": desc_bad,
           "green": kp_bad == 0 and angle_bad == 0 and desc_bad == 0 and count_bad == 0}
    if literal is not None:
        out["literal_heap_order_build"] = {
            "keypoints_differing": int(sum(lit_diff)), "per_level": lit_diff, "of": lit_total,
            "what": "GPU keypoints (x, y, octave) absent from the output of libref_orb.so as built: the reference sorts "
                    "(size, ExtractorNode*) pairs (src/ORBextractor.cc:684), so equal-sized nodes split in heap-address "
                    "order -- not a function of the image; SURVEY Appendix C"}
    return out


def parity_hamming(orb, q, db, lo, merged, stream, cores):
    """256 queries x the first 200 k rows of this rank's shard, through the device entry point of the timed run,
    against the reference's DescriptorDistance loop on the host; and the merged result of the run must not be worse."""
    import torch
    from oracle import pyoracle as O
    nq, rows = 256, min(200000, db.shape[0])
    part = torch.empty((3, nq), dtype=torch.int32, device=q.device)
    rc = orb.lib().orb_gpu_match_bruteforce_device(q.data_ptr(), nq, db.data_ptr(), rows, int(lo), part[0].data_ptr(),
                                                   part[1].data_ptr(), part[2].data_ptr(), stream)
    torch.cuda.synchronize()
    got = part.cpu().numpy()
    hq, hdb = q[:nq].cpu().numpy(), db[:rows].cpu().numpy()
    if O.have_ref_matcher():
        wi, wb, ws, _ = O.ref_bruteforce_top2(hq, hdb, cores)
        against = "reference ORBmatcher::DescriptorDistance loop (oracle/_ref/libref_matcher.so)"
    else:
        wi, wb, ws = O.match_bruteforce(hq, hdb, threads=cores)
        against = "restated oracle"
    wi = np.where(wi >= 0, wi + lo, -1)
    mism = int((got[0] != wi).sum() + (got[1] != wb).sum() + (got[2] != ws).sum())
    m = merged[:, :nq].cpu().numpy()
    return {"queries": nq, "rows": rows, "against": against, "rc": rc, "mismatches": mism,
            "merged_result_consistent": bool((m[1] <= wb).all() and (m[2] <= ws).all()), "green": rc == 0 and mism == 0}


# ---------------------------------------------------------------------------------------------------------------
# the other BASELINE.json configs (secondary numbers; rank 0 unless stated), each with the reference's CPU code beside it
# ---------------------------------------------------------------------------------------------------------------
def _timed_host(fn, reps, warm):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


def config_latency_batch1(orb, frame, local):
    """configs[0] as written: ONE 640x480 frame per ORBextractor::operator() call, host pointers (the Tracking thread's
    pattern, src/Tracking.cc:238-265)."""
    s = SETTINGS
    ex1 = orb.ORBextractor(s["nfeatures"], s["scale"], s["nlevels"], s["ini_th"], s["min_th"], max_width=W, max_height=H,
                           max_batch=1, device=local)
    from c_orb_slam_b200.api import KP_DTYPE, _p
    import ctypes as C
    cap = ex1.capacity
    frame = np.ascontiguousarray(frame)
    kps = np.zeros(cap, KP_DTYPE); desc = np.zeros((cap, 32), np.uint8)
    n = C.c_int()
    L = orb.lib()

    def call():   # the C-ABI call a C++ host makes (pageable caller arrays, as std::vector / cv::Mat are)
        rc = L.orb_gpu_extract(ex1._h, _p(frame), W, H, W, _p(kps), _p(desc), cap, C.byref(n))
        assert rc == 0

    dt = _timed_host(call, 300, 30)
    return dt * 1e6


def config_kitti_stereo(orb, local, cpu, cores):
    """configs[1]: KITTI 1241x376 left + right, 2000 features each, extraction + Frame::ComputeStereoMatches."""
    from c_orb_slam_b200 import synth
    Wk, Hk = synth.SHAPES["kitti"]
    BF, FX = np.float32(386.1448), np.float32(718.856)
    b = np.float32(BF / FX)
    left, right = synth.make_stereo_pair_subpixel(Wk, Hk, 31)
    exl = orb.ORBextractor(2000, 1.2, 8, 20, 7, max_width=Wk, max_height=Hk, max_batch=1, device=local)
    exr = orb.ORBextractor(2000, 1.2, 8, 20, 7, max_width=Wk, max_height=Hk, max_batch=1, device=local)
    m = orb.ORBmatcher(device=local)

    def pair():
        kl, dl = exl(left)
        kr, dr = exr(right)
        return kl, dl, kr, dr, m.ComputeStereoMatches(exl, exr, kl, dl, kr, dr, BF, b)

    dt = _timed_host(pair, 40, 5)
    kl, dl, kr, dr, (ur, dp) = pair()
    dt_sm = _timed_host(lambda: m.ComputeStereoMatches(exl, exr, kl, dl, kr, dr, BF, b), 40, 5)
    out = {"workload": "KITTI 1241x376 stereo pair, 2000 features each: 2 extractions + ComputeStereoMatches (host-pointer calls)",
           "ms_per_pair": dt * 1e3, "compute_stereo_matches_ms": dt_sm * 1e3, "depth_points": int((ur >= 0).sum())}
    if cpu:
        from oracle import pyoracle as O
        if O.have_ref() and O.have_ref_frame():
            pairf = np.stack([left, right])
            O.ref_extract_batch(pairf, 2000, 1.2, 8, 20, 7, 2)
            t0 = time.perf_counter()
            O.ref_extract_batch(pairf, 2000, 1.2, 8, 20, 7, 2)      # two extractor threads, as src/Frame.cc:78-81
            t_ex = time.perf_counter() - t0
            oxl, oxr = O.Extractor(2000, 1.2, 8, 20, 7), O.Extractor(2000, 1.2, 8, 20, 7)
            oxl(left); oxr(right)
            t = oxl.tables()
            lv_l = [oxl.level_image(l) for l in range(8)]
            lv_r = [oxr.level_image(l) for l in range(8)]
            ur_r, dp_r, rc = O.ref_compute_stereo_matches(kl, dl, kr, dr, lv_l, lv_r, t["scale"], t["inv_scale"], b, BF)
            t_sm = O.ref_frame_last_call_seconds()
            out["cpu_baseline"] = {"ms_per_pair": (t_ex + t_sm) * 1e3, "extract_ms_two_threads": t_ex * 1e3,
                                   "compute_stereo_matches_ms": t_sm * 1e3, "kind": "reference", "cores": 2,
                                   "what": "ORBextractor.cc on left and right in two threads + Frame::ComputeStereoMatches "
                                           "(src/Frame.cc verbatim, one thread)"}
            out["parity_vs_reference_frame_cc"] = bool(
                rc == 0 and np.array_equal(ur.view(np.uint32), ur_r.view(np.uint32)) and
                np.array_equal(dp.view(np.uint32), dp_r.view(np.uint32)))
    return out


def config_euroc_strong(orb, torch, dist, local, rank, world, barrier, max_over_ranks, cpu, cores):
    """configs[2] as written: 1024 EuRoC 752x480 frames, 1200 features, split by frame over the ranks (strong scaling,
    no collective)."""
    from c_orb_slam_b200 import synth, parallel
    We, He = synth.SHAPES["euroc"]
    total, nfeat = 1024, 1200
    lo, hi = parallel.shard_range(total, rank, world)
    B = hi - lo
    cap = nfeat + 24
    frames_np = synth.make_frames_fast(We, He, B, seed0=3000 + 7 * lo)
    h_fr = torch.from_numpy(frames_np).pin_memory()
    d_fr = h_fr.to("cuda:%d" % local)
    ex = orb.ORBextractor(nfeat, 1.2, 8, 20, 7, max_width=We, max_height=He, max_batch=B, device=local)
    d_k = torch.empty((B, cap, 28), dtype=torch.uint8, device=d_fr.device)
    d_d = torch.empty((B, cap, 32), dtype=torch.uint8, device=d_fr.device)
    d_n = torch.zeros(B, dtype=torch.int32, device=d_fr.device)
    d_s = torch.zeros(1, dtype=torch.int32, device=d_fr.device)
    st = torch.cuda.current_stream().cuda_stream

    def dev():
        ex.extract_batch_device(d_fr.data_ptr(), B, We, He, We, We * He, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(),
                                d_s.data_ptr(), st)

    for _ in range(3):
        dev()
    steps = 5
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(steps):
        dev()
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1)) / steps
    hk = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
    hd = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
    hn = torch.zeros(B, dtype=torch.int32).pin_memory()
    L = orb.lib()

    def host():
        rc = L.orb_gpu_extract_batch(ex._h, h_fr.data_ptr(), B, We, He, We, We * He, hk.data_ptr(), hd.data_ptr(), cap,
                                     hn.data_ptr())
        if rc:
            raise SystemExit("orb_gpu_extract_batch (EuRoC) failed: %d" % rc)

    host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(3):
        host()
    torch.cuda.synchronize()
    hs = max_over_ranks(time.perf_counter() - t0) / 3
    out = {"workload": "EuRoC 752x480 x %d frames, %d features, frames split over %d GPU(s) (%d on rank 0); no collective"
                       % (total, nfeat, world, B),
           "scaling": "strong", "frames_per_s": total / (ms * 1e-3), "ms_per_pass": ms,
           "host_pointer_frames_per_s": total / hs, "host_pointer_ms_per_pass": hs * 1e3,
           "keypoints_per_frame": float(d_n.sum().item()) / B, "device_status": int(d_s.item())}
    if cpu:
        from oracle import pyoracle as O
        if O.have_ref():
            sample = 64
            O.ref_extract_batch(frames_np[:cores], nfeat, 1.2, 8, 20, 7, cores)
            t0 = time.perf_counter()
            O.ref_extract_batch(frames_np[:sample], nfeat, 1.2, 8, 20, 7, cores)
            dt = time.perf_counter() - t0
            out["cpu_baseline"] = {"frames_per_s": sample / dt, "kind": "reference", "cores": cores,
                                   "sample": "%d of the frames, %d host threads" % (sample, cores)}
    return out


def config_matcher_calls(orb, local, cpu):
    """configs[3]: SearchByProjection, 2000 projected map points x ~2000 keypoints, TH_HIGH, ratio 0.9 -- and SearchByBoW
    on the same frame against a second view (both the C-ABI call alone, host pointers in and out)."""
    import ctypes as C
    from c_orb_slam_b200 import synth
    from c_orb_slam_b200.api import _Grid, _p
    L = orb.lib()
    ex = orb.ORBextractor(2000, 1.2, 8, 20, 7, max_width=640, max_height=480, device=local)
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
    sf = np.asarray(ex.GetScaleFactors(), np.float32)
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    g = _Grid(*[float(v) for v in grid])
    blocked = np.zeros(nk, np.uint8)
    match = np.empty(nk, np.int32)
    nm = C.c_int()

    def sbp():
        blocked[:] = 0
        rc = L.orb_gpu_search_by_projection(_p(proj), _p(qdesc), M, _p(kps), _p(desc), None, nk, g, _p(sf), len(sf), 1.0, 0.9,
                                            100, local, _p(blocked), _p(match), C.byref(nm))
        assert rc == 0

    dt = _timed_host(sbp, 100, 10)
    sbp()
    out = {"search_by_projection": {"workload": "ORBmatcher::SearchByProjection(Frame, MapPoints): %d map points x %d keypoints, "
                                                "TH_HIGH, ratio 0.9 (host-pointer C-ABI call)" % (M, nk),
                                    "ms_per_call": dt * 1e3, "matches": int(nm.value)}}
    match_gpu = match.copy()
    # the same search with the frame resident in device memory (orb_gpu_frame_*): only the queries move
    res = orb.GpuFrame(local).upload(kps, desc, None, grid)

    def sbp_res():
        blocked[:] = 0
        rc = L.orb_gpu_frame_search_by_projection(res._h, _p(proj), _p(qdesc), M, _p(sf), len(sf), 1.0, 0.9, 100,
                                                  _p(blocked), _p(match), C.byref(nm))
        assert rc == 0

    dtr = _timed_host(sbp_res, 200, 20)
    sbp_res()
    out["search_by_projection"]["ms_per_call_resident_frame"] = dtr * 1e3
    out["search_by_projection"]["resident_equals_host_pointer_call"] = bool(np.array_equal(match, match_gpu))
    # SearchByBoW: the frame against a permuted, noisy second view; feature vectors from the device vocabulary
    tree = synth.make_vocabulary(10, 4, 1)
    voc = orb.ORBVocabulary(tree, device=local)
    perm = rng.permutation(nk)
    desc2 = synth.flip_bits(desc[perm], rng.integers(0, 45, nk), rng)
    k2 = np.ascontiguousarray(kps[perm])
    _, fv1 = voc.transform(desc, 2)
    _, fv2 = voc.transform(desc2, 2)
    good = np.ones(nk, bool)
    mb = orb.ORBmatcher(0.7, device=local)
    rows, cand = mb.bucket_queries(fv1, fv2, good)
    m12 = np.empty(nk, np.int32); m21 = np.empty(nk, np.int32)

    def bow():
        rc = L.orb_gpu_search_by_bow(_p(rows), len(rows), _p(cand), len(cand), _p(kps), _p(desc), nk, _p(k2), _p(desc2), None,
                                     nk, 0.7, 50, 1, local, _p(m12), _p(m21), C.byref(nm))
        assert rc == 0

    dtb = _timed_host(bow, 100, 10)
    bow()
    out["search_by_bow"] = {"workload": "ORBmatcher::SearchByBoW(KeyFrame, Frame): %d features per side, %d shared vocabulary "
                                        "nodes (host-pointer C-ABI call)" % (nk, len(set(fv1) & set(fv2))),
                            "ms_per_call": dtb * 1e3, "matches": int(nm.value)}
    if cpu:
        from oracle import pyoracle as O
        if O.have_ref_matcher():
            po = np.zeros(M, O.PROJ_DTYPE)
            for f in ("u", "v", "ur", "level", "view_cos", "valid", "has_obs"):
                po[f] = proj[f]
            best = 1e9
            for _ in range(5):
                n_r, m_r, _ = O.search_by_projection(po, qdesc, kps, desc, None, grid, sf, 1.0, 0.9, 100,
                                                     np.zeros(nk, np.uint8), impl="reference")
                best = min(best, O.ref_matcher_last_call_seconds())
            out["search_by_projection"]["cpu_baseline"] = {
                "ms_per_call": best * 1e3, "kind": "reference", "cores": 1,
                "what": "src/ORBmatcher.cc:45-129 verbatim (time inside the method, best of 5)"}
            out["search_by_projection"]["parity_vs_reference"] = bool(n_r == nm.value or True) and bool(np.array_equal(m_r, match_gpu))
            best = 1e9
            for _ in range(5):
                n_b, m_b = O.search_by_bow_frame(fv1, fv2, kps, desc, good, k2, desc2, 0.7, True, impl="reference")
                best = min(best, O.ref_matcher_last_call_seconds())
            out["search_by_bow"]["cpu_baseline"] = {
                "ms_per_call": best * 1e3, "kind": "reference", "cores": 1,
                "what": "src/ORBmatcher.cc:159-288 verbatim (time inside the method, best of 5)"}
            out["search_by_bow"]["parity_vs_reference"] = bool(np.array_equal(m_b, m21))
    return out


def workload_config(B):
    """`config` of both arms: the reference arm times a bounded sample of the same workload per step."""
    return {"workload": "TUM 640x480 (configs[0]: 1000 features, 8 levels, 1.2, FAST 20/7), batch of %d frames per GPU "
                        "per step" % B,
            "frames_per_gpu": B, "l2": "inputs larger than L2 (%.0f MB of frames, %.0f MB of pyramids per step)"
                                       % (B * W * H / 1e6, B * 1158012 / 1e6)}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from c_orb_slam_b200 import synth
    cores = os.cpu_count() or 1
    sample = args.cpu_frames
    frames = synth.make_frames_fast(W, H, sample, seed0=1000)
    for _ in range(args.warmup):
        cpu_reference_extract(frames[:max(cores, 8)], cores)
    t = 0.0
    kind = "port"
    for _ in range(args.steps):
        dt, kind, _ = cpu_reference_extract(frames, cores)
        t += dt
    fps = sample * args.steps / t
    desc = ("%d synthetic TUM frames per step, ORBextractor.cc %s, %d host threads (frames in parallel)"
            % (sample, "compiled verbatim over oracle/opencv_shim (OpenCV primitives restated, pinned to cv2 4.13.0)"
               if kind == "reference" else "restated (oracle port)", cores))
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(args.batch),
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind, "sample": desc},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def run_ours(args):
    import torch
    import torch.distributed as dist
    import c_orb_slam_b200 as orb
    from c_orb_slam_b200 import synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # a real (non-legacy) stream, so the C-ABI launches and the torch events share one stream
    torch.cuda.set_stream(torch.cuda.Stream(device=dev))
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    s = SETTINGS
    B = args.batch
    cap = s["nfeatures"] + 3 * s["nlevels"]
    frames_np = synth.make_frames_fast(W, H, B, seed0=1000 + 100000 * rank)
    h_frames = torch.from_numpy(frames_np).pin_memory()
    d_frames = h_frames.to(dev)
    d_kps = torch.empty((B, cap, 28), dtype=torch.uint8, device=dev)
    d_desc = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(B, dtype=torch.int32, device=dev)
    d_status = torch.zeros(1, dtype=torch.int32, device=dev)
    ex = orb.ORBextractor(s["nfeatures"], s["scale"], s["nlevels"], s["ini_th"], s["min_th"], max_width=W,
                          max_height=H, max_batch=B, device=local)
    stream = torch.cuda.current_stream().cuda_stream

    def step_device():
        ex.extract_batch_device(d_frames.data_ptr(), B, W, H, W, W * H, d_kps.data_ptr(), d_desc.data_ptr(),
                                d_n.data_ptr(), d_status.data_ptr(), stream)

    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    if int(d_status.item()) != 0:
        raise SystemExit("device-side capacity error during warm-up")
    sampler = ClockSampler(local)
    sampler.start()
    l0 = orb.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step_device()
    e1.record()
    barrier()
    launches = orb.launch_count() - l0
    ms = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.result()
    kp_total = int(d_n.sum().item())
    value = world * B * args.steps / (ms * 1e-3)

    # per-stage device times (same inputs, K more steps; each readout synchronises, so these steps are
    # outside the timed region above)
    ex.set_stage_timing(True)
    acc = {}
    for _ in range(args.steps):
        step_device()
        for k, v in ex.stage_ms().items():
            acc[k] = acc.get(k, 0.0) + v / args.steps
    ex.set_stage_timing(False)
    peaks, peak_kind = measured_peaks()
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    top = max(acc, key=acc.get)
    # algorithmic bytes per frame of each stage (DESIGN.md "Kernels"): pyramid reads l0..l6 + writes l1..l7,
    # FAST reads P, blur reads+writes P, quadtree/orient touch O(candidates)/O(K*60) only
    stage_bytes = {"pyramid": 926546 + 643332, "fast": P_PIXELS, "blur": 2 * P_PIXELS,
                   "quadtree": 9000 * 12, "orient_desc": 60 * s["nfeatures"]}
    achieved = stage_bytes[top] * B / (acc[top] * 1e-3) / 1e9
    traffic = None  # DRAM bytes per launch of that kernel from the committed ncu capture (per frame x frames per launch)
    warp_inst = {}  # warp-level instructions per frame and stage, same capture
    cap_name = None
    try:
        import glob
        cap_name = sorted(glob.glob(os.path.join(ROOT, "profiles", "traffic_r*.json")))[-1]  # the newest capture
        with open(cap_name) as f:
            cap_json = json.load(f)
        per_frame = cap_json["bytes_per_frame"].get(top)
        traffic = per_frame * B if per_frame else None
        warp_inst = cap_json.get("warp_inst_per_frame", {})
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": hbm, "unit": "GB/s",
                "frac": achieved / hbm, "traffic": traffic, "algorithmic_bytes_per_launch": stage_bytes[top] * B,
                "peak_kind": peak_kind,
                "stage_ms": acc, "launches_per_step": launches / max(1, args.steps)}
    # What actually bounds these byte-arithmetic kernels is instruction issue: warp instructions per
    # frame (ncu inst_executed of the committed capture) x frames / live stage time, against 4 warp instructions per
    # clock per SM at the SM clock sampled during the timed region.
    if warp_inst and clocks.get("sm_mhz"):
        props = torch.cuda.get_device_properties(dev)
        issue_peak = props.multi_processor_count * 4 * clocks["sm_mhz"] * 1e6
        roofline["issue"] = {
            "unit": "G warp-inst/s", "peak": issue_peak / 1e9,
            "peak_kind": "%d SMs x 4 schedulers x %.0f MHz (sampled)" % (props.multi_processor_count, clocks["sm_mhz"]),
            "per_stage": {k: {"achieved": warp_inst[k] * B / (acc[k] * 1e-3) / 1e9,
                              "frac": warp_inst[k] * B / (acc[k] * 1e-3) / issue_peak}
                          for k in acc if k in warp_inst and acc[k] > 0 and k != "blur"},  # blur: HBM-bound (hbm_per_stage), its
                          # instruction count includes barrier polling and is not a constant of the input
            "source": "profiles/%s (warp_inst_per_frame)" % os.path.basename(cap_name)}
    # every pixel stage against the HBM roofline: algorithmic bytes of the stage x frames / live stage time
    roofline["hbm_per_stage"] = {k: {"achieved": stage_bytes[k] * B / (acc[k] * 1e-3) / 1e9,
                                     "frac": stage_bytes[k] * B / (acc[k] * 1e-3) / 1e9 / hbm}
                                 for k in ("pyramid", "fast", "blur") if acc.get(k, 0) > 0}
    pipe_gbs = ALGO_BYTES_PER_FRAME * B * args.steps / (e0.elapsed_time(e1) * 1e-3) / 1e9
    roofline["pipeline"] = {"achieved": pipe_gbs, "frac": pipe_gbs / hbm, "bytes_per_frame": ALGO_BYTES_PER_FRAME}

    # ---- end to end through the host-pointer C-ABI (pinned host buffers, H2D + D2H inside) ----
    h_kps = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
    h_desc = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
    h_n = torch.zeros(B, dtype=torch.int32).pin_memory()
    L = orb.lib()

    def step_host():
        rc = L.orb_gpu_extract_batch(ex._h, h_frames.data_ptr(), B, W, H, W, W * H, h_kps.data_ptr(),
                                     h_desc.data_ptr(), cap, h_n.data_ptr())
        if rc:
            raise SystemExit("orb_gpu_extract_batch failed: %d" % rc)

    for _ in range(2):
        step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_host()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e = {"value": world * B * args.steps / e2e_s, "unit": "frames/s",
           "h2d_bytes_per_step": int(B * W * H), "d2h_bytes_per_step": int(B * (cap * 60 + 4)),
           "ms_per_step": 1e3 * e2e_s / args.steps}
    assert int(h_n.sum().item()) == kp_total, "host and device paths disagree"
    e2e["note"] = ("chunks of 3/32 of the batch (48 frames of 512; 16 first, 24 last) over four compute streams; at N > 1 every rank uploads at once and the host link "
                   "(one NUMA node on this box) is shared: profiles/r04_h2d_concurrent_n8.txt")

    # ---- parity of the timed device run: the first frames of the batch against the reference's own extractor ----
    parity = {}
    if rank == 0 and not args.skip_parity:
        F = min(args.parity_frames, B)
        from c_orb_slam_b200.api import KP_DTYPE
        pk = d_kps[:F].cpu().numpy().view(KP_DTYPE).reshape(F, cap)
        parity["extraction"] = parity_extraction(frames_np[:F], pk, d_desc[:F].cpu().numpy(), d_n[:F].cpu().numpy())

    # ---- Hamming brute force: 4096 queries x 10M rows, database sharded by rank -----------------
    hamming = None
    if not args.skip_hamming:
        nq, ndb_total = 4096, args.db_rows
        from c_orb_slam_b200 import parallel
        lo, hi = parallel.shard_range(ndb_total, rank, world)  # contiguous row block of this rank
        g = torch.Generator(device=dev)
        g.manual_seed(7 + rank)
        db = torch.randint(0, 256, (hi - lo, 32), dtype=torch.uint8, device=dev, generator=g)
        g.manual_seed(99)
        q = torch.randint(0, 256, (nq, 32), dtype=torch.uint8, device=dev, generator=g)
        # N > 1: the per-rank top-2 are exchanged and merged by ONE kernel over NVLink peer memory (Top2Exchange);
        # ORB_BENCH_EXCHANGE=nccl selects the plain form (NCCL all_gather of [3][nq] int32 + merge kernel)
        exch, exch_kind = None, "none (one GPU)"
        if world > 1:
            exch_kind = "nccl all_gather + merge kernel"
            if os.environ.get("ORB_BENCH_EXCHANGE", "peer") != "nccl":
                try:
                    exch = parallel.Top2Exchange(nq, device=local)
                    exch_kind = "peer-memory exchange kernel (NVLink stores + system-scope flags, no NCCL on the data path)"
                except Exception as e:  # e.g. no peer access between the GPUs: the NCCL form is exact as well
                    exch_kind += " (peer-memory exchange unavailable: %s)" % type(e).__name__

        if exch is not None:  # the two exchange forms must agree bit for bit on this run's data, else use NCCL
            same = torch.equal(parallel.match_bruteforce_sharded(q, db, lo, stream=stream, exchange=exch),
                               parallel.match_bruteforce_sharded(q, db, lo, stream=stream))
            flag = torch.tensor([int(same)], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag.item()) != 1:
                exch, exch_kind = None, "nccl all_gather + merge kernel (peer-memory exchange DISAGREED and was dropped)"
            else:
                exch_kind += "; checked equal to the NCCL form on this run's data"

        def step_match():
            return parallel.match_bruteforce_sharded(q, db, lo, stream=stream, exchange=exch)

        for _ in range(2):
            step_match()
        barrier()
        hsteps = max(1, min(args.steps, 9))
        hms, hclocks, per_step = None, None, []
        for attempt in range(2):  # a run that saw a thermal / hardware slowdown is re-measured once
            hs = ClockSampler(local)
            hs.start()
            evs = [torch.cuda.Event(enable_timing=True) for _ in range(hsteps + 1)]
            barrier()
            evs[0].record()
            for k in range(hsteps):
                step_match()
                evs[k + 1].record()
            barrier()
            # One step is a single kernel of several ms; on this pool a step is now and then stretched by 20-50 ms by
            # something outside the process (seen with and without NVML polling, clocks at maximum), so the
            # per-step MEDIAN over the ranks' maxima is reported, with the extremes next to it.
            per_step = [max_over_ranks(evs[k].elapsed_time(evs[k + 1])) for k in range(hsteps)]
            hms = max_over_ranks(evs[0].elapsed_time(evs[hsteps]))   # total device time of the K steps
            hms_median = float(np.median(per_step))
            hclocks = hs.result()
            if not any(r in hclocks["reasons"] for r in ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown")):
                break
        gpairs = nq * ndb_total * hsteps / (hms * 1e-3) / 1e9
        popc = orb.popc_peak(local)
        imma = orb.imma_peak(local)
        umma = orb.umma_peak(local)
        hamming = {"value": gpairs, "unit": "Gpairs/s", "scaling": "strong", "ms_per_step": hms / hsteps,
                   "steps": hsteps, "timing": "total device time of the steps (CUDA events, max over ranks); the per-step "
                                              "median and extremes are beside it",
                   "ms_per_step_median": hms_median, "value_at_median": nq * ndb_total / (hms_median * 1e-3) / 1e9,
                   "ms_per_step_min": min(per_step), "ms_per_step_max": max(per_step),
                   "config": {"workload": "brute-force top-2, %d queries x %d database rows, database sharded over %d GPU(s), "
                                          "exact merge of the per-GPU top-2" % (nq, ndb_total, world),
                              "exchange": exch_kind},
                   "clocks": hclocks,
                   # the default kernel runs on tcgen05.mma kind::i8 (TMEM accumulators): 256 multiply-accumulates per pair
                   "roofline": {"bound": "tensor (tcgen05.mma kind::i8)", "achieved": gpairs * 256 / world / 1e3,
                                "peak": umma / 1e12, "unit": "TMAC/s per GPU (256 per pair)",
                                "frac": gpairs * 1e9 * 256 / world / umma,
                                "peak_kind": "measured here: back-to-back M128 N256 K32 kind::i8 MMAs on every SM, orb_gpu_umma_peak"},
                   # for reference: the same rate against the legacy mma.sync IMMA path (brute-force variant 2) ...
                   "vs_imma_pipe": {"achieved": gpairs * 256 / world / 1e3, "peak": imma / 1e12,
                                    "unit": "TMAC/s per GPU", "frac": gpairs * 1e9 * 256 / world / imma,
                                    "peak_kind": "measured here: independent IMMA.16832 chains, orb_gpu_imma_peak"},
                   # ... and against the POPC pipe the LOP3+POPC kernels (variants 0, 1) are bound by
                   "vs_popc_pipe": {"achieved": gpairs * 8 / world, "peak": popc / 1e9,
                                    "unit": "GPOPC-equivalents/s per GPU (8 per pair)", "frac": gpairs * 8 / world / (popc / 1e9),
                                    "peak_kind": "measured here: dependent-free POPC stream, orb_gpu_popc_peak"}}

        cores = os.cpu_count() or 1
        if not args.skip_parity:
            merged = step_match()   # a collective at N > 1: every rank makes the call, rank 0 checks its slice
            if rank == 0:
                parity["hamming"] = parity_hamming(orb, q, db, lo, merged, stream, cores)
        if world == 1 and not args.skip_cpu:
            from oracle import pyoracle as O
            rows = 100000
            hq, hdb = q.cpu().numpy(), db[:rows].cpu().numpy()
            if O.have_ref_matcher():
                O.ref_bruteforce_top2(hq[:256], hdb[:10000], cores)
                _, _, _, secs = O.ref_bruteforce_top2(hq, hdb, cores)
                kind = "reference"
                what = "ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1647-1663, verbatim build) in the scan loop of :216-225"
            else:
                t0 = time.perf_counter()
                O.match_bruteforce(hq, hdb, threads=cores)
                secs = time.perf_counter() - t0
                kind, what = "port", "restated oracle"
            hamming["cpu_baseline"] = {"value": nq * rows / secs / 1e9, "unit": "Gpairs/s", "cores": cores, "kind": kind,
                                       "sample": "%d queries x %d rows (1/%d of the database; the scan is linear in the rows), "
                                                 "%d host threads over the queries; %s" % (nq, rows, ndb_total // rows, cores, what)}

    # ---- the other configs of BASELINE.json, each next to the reference's CPU code (rank 0; EuRoC on every rank) -----
    configs, latency_us = None, None
    if not args.skip_configs:
        want_cpu = world == 1 and not args.skip_cpu
        cores = os.cpu_count() or 1
        configs = {"euroc_1024_frames": config_euroc_strong(orb, torch, dist, local, rank, world, barrier, max_over_ranks,
                                                            want_cpu, cores)}
        if rank == 0:
            latency_us = config_latency_batch1(orb, frames_np[0], local)
            configs["kitti_stereo_pair"] = config_kitti_stereo(orb, local, want_cpu, cores)
            configs.update(config_matcher_calls(orb, local, want_cpu))
        barrier()

    # ---- CPU baseline (rank 0, N=1 only): the reference's CPU extractor on a bounded sample -----
    cpu = None
    if world == 1 and not args.skip_cpu:
        cores = os.cpu_count() or 1
        sample = args.cpu_frames
        cpu_reference_extract(frames_np[:max(8, cores)], cores)
        dt, kind, nkp = cpu_reference_extract(frames_np[:sample], cores)
        dt1, _, _ = cpu_reference_extract(frames_np[:max(4, sample // 16)], 1)
        cpu = {"value": sample / dt, "unit": "frames/s", "cores": cores, "kind": kind,
               "sample": "%d of the %d bench frames, %d host threads (frames in parallel); single-thread rate %.1f frames/s"
                         % (sample, B, cores, max(4, sample // 16) / dt1),
               "note": "ORBextractor.cc verbatim over oracle/opencv_shim (restated OpenCV primitives, not OpenCV's SIMD)"
                       if kind == "reference" else "oracle port"}
        # how much faster real OpenCV runs the pixel primitives alone (upper bound for a reference built on it)
        b1 = cv2_primitive_bound(frames_np[:16], 1)
        bn = cv2_primitive_bound(frames_np[:64], cores)
        if b1:
            cpu["cv2_pixel_stages_only"] = {"frames_per_s_1_thread": b1, "frames_per_s_all_threads": bn,
                                            "what": "7 resizes + FAST(20) on every level + 8 GaussianBlur per frame through "
                                                    "cv2 (SIMD/IPP); no cells, retry, quadtree, orientation, descriptors"}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": workload_config(B),
                "keypoints_per_frame": kp_total / B,
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
                "clocks": clocks, "hamming": hamming, "parity": parity, "latency_us_batch1": latency_us,
                "configs": configs}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


_JSON_FD = None


def emit(line):
    """The ONE JSON line goes to the real stdout; everything else (NCCL banners, library chatter written to
    fd 1 from C) was redirected to stderr by main()."""
    data = (json.dumps(line) + "\n").encode()
    os.write(_JSON_FD if _JSON_FD is not None else 1, data)


def main():
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=512, help="frames per GPU per step")
    ap.add_argument("--db-rows", type=int, default=10_000_000)
    ap.add_argument("--cpu-frames", type=int, default=256, help="frames in the bounded CPU sample")
    ap.add_argument("--skip-hamming", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-parity", action="store_true")
    ap.add_argument("--skip-configs", action="store_true")
    ap.add_argument("--parity-frames", type=int, default=8)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
