"""The product's C++ drop-in class ORB_SLAM2::ORBmatcher (c_orb_slam_b200/csrc/adapter/ORBmatcher.h + ORBmatcher.cc)
next to the reference's own src/ORBmatcher.cc compiled verbatim, both under ONE harness (oracle/ref_matcher_wrapper.cc
over the stand-in Frame / KeyFrame / MapPoint of oracle/slam_shim): tests/cpp/_build/libdropin_orbmatcher.so and
oracle/_ref/libref_matcher.so export the same refm_* entry points.  Every test of tests/test_ref_matcher.py is
re-collected here with a `ref` object that repeats each call into the reference in the drop-in library on copies of
the same arrays and requires identical results: return values, mvpMapPoints / vpMatched / vpMatches12 / vnMatches12 /
vMatchedPairs as indices, the Replace / AddObservation actions of Fuse, vpReplacePoint, and the in/out arguments
(kp_blocked, vbPrevMatched).  All eleven ORBmatcher methods (include/ORBmatcher.h:48-83) run on the GPU this way."""
import copy
import ctypes as C
import os

import numpy as np
import pytest

import test_ref_matcher as T

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN = os.path.join(ROOT, "tests", "cpp", "_build", "libdropin_orbmatcher.so")

_IMPL_CALLS = ("descriptor_distance", "search_by_projection", "search_by_bow_frame", "search_by_bow_keyframes",
               "search_for_triangulation", "search_for_initialization")
_REF_CALLS = ("ref_search_last_frame", "ref_search_keyframe_points", "ref_search_sim3_projection", "ref_fuse",
              "ref_search_by_sim3", "ref_fuse_sim3")


def _same(a, b, where):
    if isinstance(a, np.ndarray):
        if a.dtype.names and "radius" in a.dtype.names:
            return          # the windows the reference recorded from its GetFeaturesInArea calls; the drop-in makes none
        assert a.dtype == b.dtype and a.shape == b.shape, where
        if a.dtype.names:           # field by field: the padding bytes of a record are not data
            for f in a.dtype.names:
                assert a[f].tobytes() == b[f].tobytes(), "%s.%s" % (where, f)
        else:
            assert a.tobytes() == b.tobytes(), where
    elif isinstance(a, (tuple, list)):
        assert len(a) == len(b), where
        for i, (x, y) in enumerate(zip(a, b)):
            _same(x, y, "%s[%d]" % (where, i))
    elif isinstance(a, dict):
        for k in a:
            _same(a[k], b[k], "%s[%r]" % (where, k))
    else:
        assert a == b, where


class _Both:
    """The oracle module, except that calls into the verbatim reference matcher are also made into the drop-in."""
    calls = 0

    def __init__(self, oracle):
        self._o = oracle

    def __getattr__(self, name):
        f = getattr(self._o, name)
        if name not in _IMPL_CALLS and name not in _REF_CALLS:
            return f

        def both(*a, **k):
            if name in _IMPL_CALLS and k.get("impl") != "reference":
                return f(*a, **k)
            a2, k2 = copy.deepcopy(a), copy.deepcopy(k)
            want = f(*a, **k)
            with self._o.matcher_library(DROPIN):
                got = f(*a2, **k2)
            _same(want, got, name + " result")
            _same(a, a2, name + " in/out argument")
            _Both.calls += 1
            return want
        return both


@pytest.fixture(scope="module")
def ref(oracle, gpu):
    if not oracle.have_ref_matcher() or not os.path.exists(DROPIN):
        pytest.skip("libref_matcher.so / libdropin_orbmatcher.so not built (tests/cpp/Makefile, needs /root/reference)")
    return _Both(oracle)


# every case of the CPU file, now with the drop-in class in the loop
from test_ref_matcher import (test_descriptor_distance_and_constants, test_search_by_projection_equals_reference,  # noqa: E402,F401
                              test_search_by_bow_frame_equals_reference, test_search_by_bow_keyframes_equals_reference,
                              test_search_for_triangulation_equals_reference,
                              test_search_for_initialization_equals_reference, test_search_last_frame_equals_reference,
                              test_search_keyframe_points_equals_reference, test_search_sim3_projection_equals_reference,
                              test_fuse_equals_reference, test_search_by_sim3_equals_reference,
                              test_fuse_sim3_equals_reference)


def test_dropin_class_constants_and_device(ref, gpu):
    """TH_HIGH / TH_LOW / HISTO_LENGTH of the drop-in class, and that the calls above really ran in it on the GPU."""
    L = C.CDLL(DROPIN)
    th = [C.c_int() for _ in range(3)]
    L.refm_constants(*[C.byref(t) for t in th])
    assert [t.value for t in th] == [100, 50, 30]
    before = gpu.launch_count()
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, 32, dtype=np.uint8); b = rng.integers(0, 256, 32, dtype=np.uint8)
    assert ref.descriptor_distance(a, b, impl="reference") == int(np.unpackbits(a ^ b).sum())
    assert gpu.launch_count() == before          # one pair is host work in the drop-in too
    assert _Both.calls > 0
