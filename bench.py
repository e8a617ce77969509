#!/usr/bin/env python
"""Benchmark of the ORB front-end hot path (BASELINE.json metric: "ORB extract frames/s (TUM 640x480,
1000 kp); Hamming Gpairs/s").

  python bench.py --gpus N --steps K --warmup W            # this repo (CUDA, sm_100a)
  python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU code on host cores

One step = ORBextractor::operator() over one batch of synthetic TUM-shaped frames (TUM1.yaml
settings: 1000 features, 8 levels, 1.2, iniThFAST 20, minThFAST 7).  `value` is frames/s with the
frames resident in HBM (device C-ABI, CUDA events on the launching stream); `e2e` is the same work
through the host-pointer C-ABI call with pinned host buffers, H2D/D2H inside the timed region.
Frames shard across ranks (independent units, no collective): weak scaling.  A secondary object
`hamming` reports the brute-force 4096 x 10M top-2 matcher (database sharded across ranks, one NCCL
all-gather of the per-rank top-2, exact merge).  Prints ONE JSON line on rank 0.
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
            "config": {"workload": "TUM 640x480, 1000 features, 8 levels, 1.2, FAST 20/7 (configs[0]); CPU sample of %d frames/step" % sample},
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
    try:
        with open(os.path.join(ROOT, "profiles", "traffic_r05.json")) as f:
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
                          for k in acc if k in warp_inst and acc[k] > 0},
            "source": "profiles/traffic_r05.json (warp_inst_per_frame)"}
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
    # the copies alone (same bytes, same pinned buffers, H2D and D2H on two streams, every rank at once): what the
    # host link leaves for any implementation of the host-pointer call at this N
    d_in = torch.empty_like(h_frames, device=dev)
    d_ok, d_od = torch.empty_like(h_kps, device=dev), torch.empty_like(h_desc, device=dev)
    s_out = torch.cuda.Stream(device=dev)

    def copies():
        d_in.copy_(h_frames, non_blocking=True)
        with torch.cuda.stream(s_out):
            h_kps.copy_(d_ok, non_blocking=True)
            h_desc.copy_(d_od, non_blocking=True)

    copies()
    barrier()
    t0 = time.perf_counter()
    for _ in range(5):
        copies()
    torch.cuda.synchronize()
    copy_s = max_over_ranks(time.perf_counter() - t0) / 5
    e2e["copies_only_ms_per_step"] = 1e3 * copy_s
    e2e["copies_only_note"] = ("pinned H2D of the frames + D2H of the outputs alone, all %d rank(s) at once: the host link's "
                               "bound on this path (%.1f GB/s H2D per GPU)" % (world, B * W * H / copy_s / 1e9))
    del d_in, d_ok, d_od

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
            hms = float(np.median(per_step)) * hsteps
            hclocks = hs.result()
            if not any(r in hclocks["reasons"] for r in ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown")):
                break
        gpairs = nq * ndb_total * hsteps / (hms * 1e-3) / 1e9
        popc = orb.popc_peak(local)
        imma = orb.imma_peak(local)
        umma = orb.umma_peak(local)
        hamming = {"value": gpairs, "unit": "Gpairs/s", "scaling": "strong", "ms_per_step": hms / hsteps,
                   "steps": hsteps, "timing": "median of per-step device times (CUDA events, max over ranks)",
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
                "config": {"workload": "TUM 640x480 (configs[0]: 1000 features, 8 levels, 1.2, FAST 20/7), "
                                       "batch of %d frames per GPU per step" % B,
                           "frames_per_gpu": B, "l2": "inputs larger than L2 (%.0f MB of frames, %.0f MB of pyramids per step)"
                                                      % (B * W * H / 1e6, B * 1158012 / 1e6),
                           "keypoints_per_frame": kp_total / B},
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
                "clocks": clocks, "hamming": hamming}
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
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
