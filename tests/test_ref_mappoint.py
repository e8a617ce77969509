"""The restated MapPoint::ComputeDistinctiveDescriptors (row N4) against the reference's own src/MapPoint.cc compiled
VERBATIM with its own include/MapPoint.h (oracle/_ref/libref_mappoint.so, oracle/Makefile target `refmappoint`)."""
import numpy as np
import pytest

from c_orb_slam_b200 import synth


def test_distinctive_descriptors_equal_reference(oracle):
    if not oracle.have_ref_mappoint():
        pytest.skip("oracle/_ref/libref_mappoint.so not built (needs /root/reference)")
    rng = np.random.default_rng(4)
    lists = []
    for p in range(300):
        n = int(rng.integers(1, 40)) if p % 7 else int(rng.integers(1, 4))
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        d = synth.flip_bits(np.repeat(base[None], n, 0), rng.integers(0, 50, n), rng)
        if p % 5 == 0 and n > 2:            # exact duplicates: equal medians, the first one must win (strict '<', :296)
            d[n // 2] = d[0]
            d[-1] = d[1]
        lists.append(d)
    best_o = oracle.distinctive_descriptors(lists)
    best_r = oracle.ref_distinctive_descriptors(lists)
    # the verbatim call reports the first observation whose descriptor equals the one kept; the oracle reports BestIdx
    for d, bo, br in zip(lists, best_o, best_r):
        assert br >= 0 and np.array_equal(d[bo], d[br]) and br <= bo
        assert bo == br or np.array_equal(d[bo], d[br])
    assert (best_o == best_r).mean() > 0.9
    # observations from bad key frames are left out (:263-267): same as the oracle on the filtered list
    bad = [rng.random(len(d)) < 0.3 for d in lists]
    best_rb = oracle.ref_distinctive_descriptors(lists, bad)
    for d, b, br in zip(lists, bad, best_rb):
        keep = np.flatnonzero(~b)
        if len(keep) == 0:
            assert br == -1
            continue
        bo = keep[oracle.distinctive_descriptors([d[keep]])[0]]
        assert br >= 0 and np.array_equal(d[bo], d[br])
