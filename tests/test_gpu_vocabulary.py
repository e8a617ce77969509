"""N3 parity: DBoW2 transform (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1259) on a synthetic vocabulary
tree (ORBvoc.txt is not part of the reference checkout): per-feature word / weight / node bit-exact, and the
BowVector (float64, TF_IDF + L1) and FeatureVector built from them equal to the oracle's std::map restatement."""
import numpy as np
import pytest

from c_orb_slam_b200 import synth

pytestmark = pytest.mark.gpu


def _features(tree, n, seed, max_flips=60):
    rng = np.random.default_rng(seed)
    leaf = np.flatnonzero(tree["word_id"] >= 0)
    f = synth.flip_bits(tree["desc"][rng.choice(leaf, n)], rng.integers(0, max_flips, n), rng)
    f[: n // 10] = rng.integers(0, 256, (n // 10, 32), dtype=np.uint8)   # far from every word: many near ties
    return f


@pytest.mark.parametrize("k,depth,levelsup,seed", [(10, 4, 2, 1), (10, 5, 4, 2), (7, 3, 4, 3), (33, 2, 1, 4)])
def test_vocabulary_transform(gpu, oracle, k, depth, levelsup, seed):
    tree = synth.make_vocabulary(k, depth, seed)
    voc = gpu.ORBVocabulary(tree)
    f = _features(tree, 2000, seed)
    (w_o, v_o, n_o), (bid_o, bval_o), fv_o = oracle.voc_transform(tree, f, levelsup)
    w_g, v_g, n_g = voc.transform_features(f, levelsup)
    assert np.array_equal(w_g, w_o) and np.array_equal(n_g, n_o)
    assert np.array_equal(v_g.view(np.uint64), v_o.view(np.uint64))
    (bid_g, bval_g), fv_g = voc.transform(f, levelsup)
    assert np.array_equal(bid_g, bid_o) and np.array_equal(bval_g.view(np.uint64), bval_o.view(np.uint64))
    assert fv_g == fv_o
    assert (v_o == 0).any() or depth < 3            # stopped words exist and are left out of both maps
    if depth - levelsup <= 0:
        assert set(fv_g) == {0}
    assert len(bid_g) > 100 and abs(bval_g.sum() - 1.0) < 1e-12


def test_vocabulary_feeds_search_by_bow(gpu, oracle):
    """Frame::ComputeBoW -> SearchByBoW chain: feature vectors from the device transform drive the bucket search."""
    tree = synth.make_vocabulary(10, 4, 7)
    voc = gpu.ORBVocabulary(tree)
    rng = np.random.default_rng(7)
    leaf = np.flatnonzero(tree["word_id"] >= 0)
    d1 = synth.flip_bits(tree["desc"][rng.choice(leaf, 1500)], rng.integers(0, 30, 1500), rng)
    perm = rng.permutation(1500)
    d2 = synth.flip_bits(d1[perm], rng.integers(0, 45, 1500), rng)
    k = np.zeros(1500, gpu.KP_DTYPE)
    _, fv1 = voc.transform(d1, 2)
    _, fv2 = voc.transform(d2, 2)
    _, _, fvo1 = oracle.voc_transform(tree, d1, 2)
    _, _, fvo2 = oracle.voc_transform(tree, d2, 2)
    assert fv1 == fvo1 and fv2 == fvo2
    good = np.ones(1500, bool)
    n_o, m21_o = oracle.search_by_bow_frame(fvo1, fvo2, k, d1, good, k, d2, 0.7, False)
    n_g, m12, m21 = gpu.ORBmatcher(0.7, checkOri=False).SearchByBoW(
        dict(kps=k, desc=d1, feat_vec=fv1), dict(kps=k, desc=d2, feat_vec=fv2), good)
    assert n_g == n_o and np.array_equal(m21, m21_o) and n_g > 500
    right = (perm[np.flatnonzero(m21 >= 0)] == m21[m21 >= 0]).mean()
    assert right > 0.95


def test_vocabulary_argument_checks(gpu):
    tree = synth.make_vocabulary(4, 2, 1)
    bad = dict(tree)
    bad["children"] = tree["children"].copy()
    bad["children"][0] = 0                               # the root as somebody's child: a cycle
    with pytest.raises(gpu.OrbGpuError):
        gpu.ORBVocabulary(bad)
    voc = gpu.ORBVocabulary(tree)
    w, v, n = voc.transform_features(np.zeros((0, 32), np.uint8))
    assert len(w) == 0
