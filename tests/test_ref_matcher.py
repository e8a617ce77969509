"""The restated matcher oracle (oracle/orb_oracle.cc) against the reference's own src/ORBmatcher.cc compiled
VERBATIM over oracle/slam_shim (oracle/_ref/libref_matcher.so, oracle/Makefile target `refmatch`), on the same
randomised cases the GPU parity tests use.  CPU only; skipped where the reference build is absent."""
import numpy as np
import pytest

from c_orb_slam_b200 import synth
from test_gpu_bow_init import _two_views
from test_gpu_matcher import _projection_case


@pytest.fixture(scope="module")
def ref(oracle):
    if not oracle.have_ref_matcher():
        pytest.skip("oracle/_ref/libref_matcher.so not built (needs /root/reference)")
    oracle.ref_matcher()
    return oracle


def test_descriptor_distance_and_constants(ref):
    import ctypes as C
    th = [C.c_int() for _ in range(3)]
    ref.ref_matcher().refm_constants(*[C.byref(t) for t in th])
    assert [t.value for t in th] == [100, 50, 30]     # TH_HIGH, TH_LOW, HISTO_LENGTH (src/ORBmatcher.cc:37-39)
    rng = np.random.default_rng(0)
    d = rng.integers(0, 256, (400, 32), dtype=np.uint8)
    d[0] = 0; d[1] = 255
    for i in range(0, 400, 2):
        a, b = d[i], d[i + 1]
        want = int(np.unpackbits(a ^ b).sum())
        assert ref.descriptor_distance(a, b, impl="reference") == want == ref.descriptor_distance(a, b)


@pytest.mark.parametrize("th,stereo,seed", [(1.0, False, 1), (3.0, False, 2), (5.0, True, 3), (15.0, True, 4)])
@pytest.mark.parametrize("nnratio", [0.9, 0.6])
def test_search_by_projection_equals_reference(ref, th, stereo, seed, nnratio):
    kps, desc, proj, qdesc, u_right, blocked, grid, sf = _projection_case(ref, seed, th, stereo)
    n_o, m_o, b_o = ref.search_by_projection(proj, qdesc, kps, desc, u_right, grid, sf, th, nnratio, 100, blocked.copy())
    n_r, m_r, b_r = ref.search_by_projection(proj, qdesc, kps, desc, u_right, grid, sf, th, nnratio, 100, blocked.copy(),
                                             impl="reference")
    assert n_r > 100
    assert n_o == n_r and np.array_equal(m_o, m_r) and np.array_equal(b_o, b_r)


@pytest.mark.parametrize("check_ori,ratio,seed", [(True, 0.7, 1), (False, 0.9, 2), (True, 0.75, 3)])
def test_search_by_bow_frame_equals_reference(ref, check_ori, ratio, seed):
    _, rng, (k1, d1, fv1), (k2, d2, fv2) = _two_views(ref, seed)
    good1 = rng.random(len(k1)) < 0.8
    n_o, m_o = ref.search_by_bow_frame(fv1, fv2, k1, d1, good1, k2, d2, ratio, check_ori)
    n_r, m_r = ref.search_by_bow_frame(fv1, fv2, k1, d1, good1, k2, d2, ratio, check_ori, impl="reference")
    assert n_r > 200 and n_o == n_r and np.array_equal(m_o, m_r)


@pytest.mark.parametrize("check_ori,ratio,seed", [(True, 0.75, 4), (False, 0.8, 5)])
def test_search_by_bow_keyframes_equals_reference(ref, check_ori, ratio, seed):
    _, rng, (k1, d1, fv1), (k2, d2, fv2) = _two_views(ref, seed)
    good1 = rng.random(len(k1)) < 0.8
    good2 = rng.random(len(k2)) < 0.8
    n_o, m_o = ref.search_by_bow_keyframes(fv1, fv2, k1, d1, good1, k2, d2, good2, ratio, check_ori)
    n_r, m_r = ref.search_by_bow_keyframes(fv1, fv2, k1, d1, good1, k2, d2, good2, ratio, check_ori, impl="reference")
    assert n_r > 150 and n_o == n_r and np.array_equal(m_o, m_r)


@pytest.mark.parametrize("stereo,only_stereo,check_ori,seed", [
    (False, False, True, 6), (True, False, True, 7), (True, True, False, 8)])
def test_search_for_triangulation_equals_reference(ref, stereo, only_stereo, check_ori, seed):
    ex, rng, (k1, d1, fv1), (k2, d2, fv2) = _two_views(ref, seed)
    n1, n2 = len(k1), len(k2)
    has1 = rng.random(n1) < 0.3
    has2 = rng.random(n2) < 0.3
    ur1 = ur2 = None
    if stereo:
        ur1 = np.where(rng.random(n1) < 0.5, k1["x"] - 10, -1).astype(np.float32)
        ur2 = np.where(rng.random(n2) < 0.5, k2["x"] - 10, -1).astype(np.float32)
    F12 = (np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32) + rng.normal(0, 2e-6, (3, 3)).astype(np.float32))
    sf = ex.tables()["scale"].astype(np.float32)
    sigma2 = (sf * sf).astype(np.float32)
    epi = (np.float32(k2["x"][5]), np.float32(k2["y"][5]))
    args = (fv1, fv2, k1, d1, ur1, has1, k2, d2, ur2, has2, F12, epi[0], epi[1], sf, sigma2, only_stereo, check_ori)
    n_o, m_o = ref.search_for_triangulation(*args)
    n_r, m_r = ref.search_for_triangulation(*args, impl="reference")
    assert n_r > (40 if only_stereo else 150) and n_o == n_r and np.array_equal(m_o, m_r)


@pytest.mark.parametrize("check_ori,window,seed", [(True, 100, 9), (False, 100, 10), (True, 30, 11)])
def test_search_for_initialization_equals_reference(ref, check_ori, window, seed):
    _, rng, (k1, d1, _), (k2, d2, _) = _two_views(ref, seed, dup=400)
    grid = (0.0, 0.0, np.float32(64) / np.float32(640), np.float32(48) / np.float32(480))
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    for _ in range(2):     # second round from the positions the reference wrote back (:514-517)
        n_o, m_o = ref.search_for_initialization(k1, d1, prev.copy(), k2, d2, grid, window, 0.9, check_ori)
        n_r, m_r = ref.search_for_initialization(k1, d1, prev, k2, d2, grid, window, 0.9, check_ori, impl="reference")
        assert n_r > 100 and n_o == n_r and np.array_equal(m_o, m_r)
        hit = m_r >= 0
        assert np.array_equal(prev[hit, 0], k2["x"][m_r[hit]]) and np.array_equal(prev[hit, 1], k2["y"][m_r[hit]])
