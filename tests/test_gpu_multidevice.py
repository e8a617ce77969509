"""Handles on a second GPU of the same process (needs 2 GPUs; skipped otherwise).  The per-device state of the
library -- shared-memory opt-ins of the kernels, the cached memory pool, TMA tensor maps, the static scratch of the
no-alloc matcher -- must work on device 1 after device 0 was used, and from two threads at once."""
import threading

import numpy as np
import pytest

from c_orb_slam_b200 import synth

pytestmark = pytest.mark.gpu


def _two_gpus():
    import torch
    return torch.cuda.device_count() >= 2


@pytest.mark.skipif(not _two_gpus(), reason="needs two GPUs in one process")
def test_extractors_on_two_devices_of_one_process(gpu, oracle):
    W, H = synth.SHAPES["tum"]
    frames = synth.make_frames(W, H, 3, seed0=500)
    ex0 = gpu.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=3, device=0)
    n0, k0, d0 = ex0.extract_batch(frames)
    ex1 = gpu.ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=3, device=1)
    n1, k1, d1 = ex1.extract_batch(frames)
    ex_o = oracle.Extractor(1000, 1.2, 8, 20, 7)
    for f in range(3):
        ko, do = ex_o(frames[f])
        for n, k, d in ((n0, k0, d0), (n1, k1, d1)):
            assert n[f] == len(ko) and np.array_equal(k[f, :n[f]], ko) and np.array_equal(d[f, :n[f]], do)
    # both at once, one host thread per device (the SLAM threads' pattern)
    out = {}

    def run(ex, key):
        out[key] = [ex.extract_batch(frames) for _ in range(5)][-1]

    ts = [threading.Thread(target=run, args=(ex0, 0)), threading.Thread(target=run, args=(ex1, 1))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    for key in (0, 1):
        n, k, d = out[key]
        assert np.array_equal(n, n0) and np.array_equal(k, k0) and np.array_equal(d, d0)


@pytest.mark.skipif(not _two_gpus(), reason="needs two GPUs in one process")
def test_bruteforce_on_device_1(gpu, oracle):
    import torch
    q, db = synth.make_match_set(300, 30000, seed=9)
    oi, ob, os_ = oracle.match_bruteforce(q, db, threads=4)
    for dev in (0, 1, 0):
        with torch.cuda.device(dev):
            dq = torch.from_numpy(q).cuda(dev)
            ddb = torch.from_numpy(db).cuda(dev)
            out = torch.empty((3, len(q)), dtype=torch.int32, device="cuda:%d" % dev)
            rc = gpu.lib().orb_gpu_match_bruteforce_device(dq.data_ptr(), len(q), ddb.data_ptr(), len(db), 0,
                                                           out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), None)
            assert rc == 0
            torch.cuda.synchronize(dev)
            o = out.cpu().numpy()
            assert np.array_equal(o[0], oi) and np.array_equal(o[1], ob) and np.array_equal(o[2], os_)
