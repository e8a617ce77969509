"""The restated vocabulary transform (row N3) against DBoW2 as the reference vendors it, compiled VERBATIM
(oracle/_ref/libref_vocabulary.so, oracle/Makefile target `refvoc`): a synthetic tree is written in the ORBvoc.txt
text format, loaded by TemplatedVocabulary::loadFromTextFile and queried with transform(features, BowVector,
FeatureVector, levelsup) exactly as Frame::ComputeBoW does (src/Frame.cc:395-402).  CPU only."""
import numpy as np
import pytest

from c_orb_slam_b200 import synth


@pytest.fixture(scope="module")
def ref(oracle):
    if not oracle.have_ref_vocabulary():
        pytest.skip("oracle/_ref/libref_vocabulary.so not built (needs /root/reference)")
    oracle.ref_vocabulary_lib()
    return oracle


def _features(tree, n, seed, max_flips=60):
    rng = np.random.default_rng(seed)
    leaf = np.flatnonzero(tree["word_id"] >= 0)
    f = synth.flip_bits(tree["desc"][rng.choice(leaf, n)], rng.integers(0, max_flips, n), rng)
    f[: n // 10] = rng.integers(0, 256, (n // 10, 32), dtype=np.uint8)   # far from every word: many near ties
    return f


@pytest.mark.parametrize("k,depth,levelsup,seed", [(10, 4, 2, 1), (10, 5, 4, 2), (7, 3, 4, 3), (20, 2, 1, 4), (10, 4, 0, 5)])
def test_transform_equals_dbow2(ref, tmp_path, k, depth, levelsup, seed):
    tree = ref.tree_as_loaded(synth.make_vocabulary(k, depth, seed))
    path = str(tmp_path / "voc.txt")
    ref.write_vocabulary_text(tree, k, path)
    voc = ref.RefVocabulary(path)
    assert voc.words() == int((tree["word_id"] >= 0).sum())
    f = _features(tree, 2000, seed)
    _, (bid_o, bval_o), fv_o = ref.voc_transform(tree, f, levelsup)
    (bid_r, bval_r), fv_r = voc.transform(f, levelsup)
    assert len(bid_r) > 100
    assert np.array_equal(bid_o, bid_r) and np.array_equal(bval_o.view(np.uint64), bval_r.view(np.uint64))
    assert fv_o == fv_r
    if depth - levelsup <= 0:
        assert set(fv_r) == {0}
