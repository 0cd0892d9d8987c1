"""DBoW2 transform: the restated oracle (oracle/orbref_bow.cpp) against the reference's own DBoW2 compiled
unmodified (oracle/_ref/libdbow2_ref.so), and against committed golden vectors generated from it
(tests/golden/bow_reference.npz, tools/make_golden_bow.py)."""
import os

import numpy as np
import pytest

from ar_orbslam2_b200.synth import synth_bow_descriptors, synth_vocabulary, write_vocabulary_text
from oracle import orbref

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "bow_reference.npz")

# (k, L, seed, scoring, weighting): every weighting type, both norms and the no-normalisation scoring
CASES = [(10, 3, 1, 0, 0), (10, 4, 2, 0, 0), (4, 5, 3, 1, 1), (7, 3, 4, 5, 0), (9, 3, 5, 2, 2), (3, 6, 6, 5, 3), (16, 2, 7, 0, 0),
         (20, 2, 8, 1, 0), (5, 4, 9, 5, 1)]


def _same(a, b):
    assert len(a) == len(b)
    for x, y in zip(a, b):
        assert x.dtype == y.dtype and x.tobytes() == y.tobytes()


def bow_case(k, L, seed, n=1500, early=0.0):
    voc = synth_vocabulary(k, L, seed=seed, early_leaf_fraction=early)
    return voc, synth_bow_descriptors(voc, n, seed)


needs_ref = pytest.mark.skipif(not orbref.ref_dbow_available(), reason="oracle/_ref/libdbow2_ref.so not built and no reference checkout")


@needs_ref
@pytest.mark.parametrize("k,L,seed,scoring,weighting", CASES)
def test_oracle_equals_reference_dbow2(tmp_path, k, L, seed, scoring, weighting):
    voc, d = bow_case(k, L, seed)
    path = tmp_path / "voc.txt"
    write_vocabulary_text(path, voc, scoring, weighting, trailing_newline=False)
    R, O = orbref.ReferenceVocabulary(path), orbref.RefVocabulary(path=path)
    assert R.info() == O.info()
    _same(R.nodes(), O.nodes())                                   # loadFromTextFile: same tree
    for levelsup in range(0, L + 2):
        _same(R.transform_features(d, levelsup), O.transform_features(d, levelsup))
        _same(R.transform(d, levelsup), O.transform(d, levelsup))  # BowVector doubles bit-for-bit, FeatureVector CSR
    assert len(R.transform(d[:0], 4)[0]) == 0 and len(O.transform(d[:0], 4)[0]) == 0


def test_oracle_loader_ignores_the_trailing_newline(tmp_path):
    voc, d = bow_case(6, 3, 11, n=300)
    a, b = tmp_path / "a.txt", tmp_path / "b.txt"
    write_vocabulary_text(a, voc, trailing_newline=False)
    write_vocabulary_text(b, voc, trailing_newline=True)
    A, B = orbref.RefVocabulary(path=a), orbref.RefVocabulary(path=b)
    assert A.info() == B.info()
    _same(A.nodes(), B.nodes())
    _same(A.transform(d), B.transform(d))
    # the flat-array constructor describes the same tree
    F = orbref.RefVocabulary(arrays=voc)
    _same(A.transform(d, 2), F.transform(d, 2))


def test_oracle_rejects_malformed_files(tmp_path):
    p = tmp_path / "bad.txt"
    p.write_text("30 3 0 0\n0 1 " + " ".join(["0"] * 32) + " 1.0\n")     # k out of range (ref :1360)
    with pytest.raises(RuntimeError):
        orbref.RefVocabulary(path=p)
    p.write_text("10 3 0 0\nxyz\n0 1 " + " ".join(["0"] * 32) + " 1.0\n")
    with pytest.raises(RuntimeError):
        orbref.RefVocabulary(path=p)


def test_early_leaves_and_stop_words():
    voc, d = bow_case(5, 4, 21, n=800, early=0.3)
    O = orbref.RefVocabulary(arrays=voc)
    word, node, w = O.transform_features(d, 0)
    assert (w == 0).any() and (w > 0).any()
    bw, bv, fn, fo, ff = O.transform(d, 1)
    assert len(ff) == int((w > 0).sum()) and fo[-1] == len(ff)
    assert np.all(np.diff(bw) > 0) and np.all(np.diff(fn) > 0)
    assert abs(bv.sum() - 1.0) < 1e-12                                     # L1-normalised


def test_oracle_equals_golden():
    """The committed vectors were produced by the reference's compiled DBoW2 (tools/make_golden_bow.py); they pin the
    oracle where /root/reference is absent (the GPU box)."""
    g = np.load(GOLDEN)
    for c in range(int(g["ncases"])):
        k, L, seed, scoring, weighting, levelsup, n = (int(x) for x in g[f"case{c}"])
        voc, d = bow_case(k, L, seed, n=n)
        O = orbref.RefVocabulary(arrays=voc, scoring=scoring, weighting=weighting)
        got = O.transform(d, levelsup)
        for name, arr in zip(("bow_word", "bow_value", "fv_node", "fv_off", "fv_feat"), got):
            assert arr.tobytes() == g[f"c{c}_{name}"].tobytes(), (c, name)
        word, node, w = O.transform_features(d, levelsup)
        assert np.array_equal(word, g[f"c{c}_word"]) and np.array_equal(node, g[f"c{c}_node"])
