"""orb_gpu_match_bruteforce_noalloc_device (no counterpart in the reference):
the brute-force top-2 search with its scratch in a static device array instead of an allocation.  Called
through the C ABI and compared bit for bit with the existing entry point orb_gpu_match_bruteforce_device on two calls
in a row (the second smaller, so stale scratch of the first must not leak), plus the size limit.  Collected last."""
import numpy as np
import pytest

from c_orb_slam_b200 import synth

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 1 << 20   # ORB_GPU_NOALLOC_SCRATCH_BYTES, include/orb_gpu.h


def _both(L, torch, q, db, base):
    dq, ddb = torch.from_numpy(q).cuda(), torch.from_numpy(db).cuda()
    st = torch.cuda.current_stream().cuda_stream
    outs = []
    for fn in (L.orb_gpu_match_bruteforce_device, L.orb_gpu_match_bruteforce_noalloc_device):
        o = torch.full((3, len(q)), -7, dtype=torch.int32, device="cuda")
        assert fn(dq.data_ptr(), len(q), ddb.data_ptr(), len(db), base, o[0].data_ptr(), o[1].data_ptr(), o[2].data_ptr(), st) == 0
        torch.cuda.synchronize()
        outs.append(o.cpu().numpy())
    return outs


def test_noalloc_bruteforce_equals_existing_entry_point(gpu, oracle):
    import torch
    L = gpu.lib()
    L.orb_gpu_set_bruteforce_variant(3, 2)
    # call 1: 1000 queries x 150 k rows (4 query blocks -> 37 slices of scratch on 148 SMs); call 2: smaller, ragged
    for nq, ndb, base, seed in ((1000, 150001, 0, 61), (300, 40003, 1000, 62)):
        q, db = synth.make_match_set(nq, ndb, seed=seed)
        want, got = _both(L, torch, q, db, base)
        assert np.array_equal(got, want)            # bit-exact: integer distances and indices
        oi, ob, os_ = oracle.match_bruteforce(q, db, threads=8)
        assert np.array_equal(got[1], ob) and np.array_equal(got[2], os_) and np.array_equal(got[0], np.where(oi >= 0, oi + base, -1))
    # the same inputs again give the same answer (the array keeps the last call's contents; what matters is cleared)
    want, got = _both(L, torch, q, db, base)
    assert np.array_equal(got, want)


def test_noalloc_bruteforce_rejects_sizes_beyond_the_static_scratch(gpu):
    import torch
    L = gpu.lib()
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    q = torch.zeros((256, 32), dtype=torch.uint8, device="cuda")
    db = torch.zeros((256, 32), dtype=torch.uint8, device="cuda")     # never read: the call is refused before any launch
    o = torch.zeros((3, 256), dtype=torch.int32, device="cuda")
    ndb = 5 * sms * (1 << 22)                                          # five waves of slices: 5 * SMs * 256 * 12 bytes
    assert 5 * sms * 256 * 12 > SCRATCH_BYTES
    before = gpu.launch_count()
    rc = L.orb_gpu_match_bruteforce_noalloc_device(q.data_ptr(), 256, db.data_ptr(), ndb, 0, o[0].data_ptr(),
                                                   o[1].data_ptr(), o[2].data_ptr(), None)
    assert rc == -2                                                    # ORB_GPU_ERR_CAPACITY
    assert gpu.launch_count() == before
    msg = L.orb_gpu_last_error().decode()
    assert "scratch" in msg and str(SCRATCH_BYTES) in msg
    assert L.orb_gpu_match_bruteforce_noalloc_device(None, 4, db.data_ptr(), 10, 0, o[0].data_ptr(), o[1].data_ptr(),
                                                     o[2].data_ptr(), None) == -1   # ORB_GPU_ERR_ARG
