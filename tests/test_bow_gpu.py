"""GPU parity of the DBoW2 transform (orb_vocab_*) through the C ABI: against the oracle, against the
reference-generated golden vectors and, where oracle/_ref/libdbow2_ref.so travelled, against the reference's DBoW2."""
import os

import numpy as np
import pytest

from ar_orbslam2_b200.synth import synth_bow_descriptors, synth_vocabulary, write_vocabulary_text
from ar_orbslam2_b200.vocabulary import ORBVocabulary
from oracle import orbref
from test_bow_cpu import CASES, GOLDEN, _same, bow_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k,L,seed,scoring,weighting", CASES)
def test_gpu_equals_oracle_and_reference(tmp_path, k, L, seed, scoring, weighting):
    voc, d = bow_case(k, L, seed)
    path = tmp_path / "voc.txt"
    write_vocabulary_text(path, voc, scoring, weighting, trailing_newline=True)   # as saveToTextFile writes it
    G = ORBVocabulary()
    assert G.loadFromTextFile(path)
    O = orbref.RefVocabulary(path=path)
    assert G.info() == O.info()
    _same(G.nodes(), O.nodes())
    R = None
    if os.path.exists(orbref.REF_DBOW_PATH):
        p2 = tmp_path / "voc_nonl.txt"
        write_vocabulary_text(p2, voc, scoring, weighting, trailing_newline=False)
        R = orbref.ReferenceVocabulary(p2)
    for levelsup in range(0, L + 2):
        _same(G.transform_features(d, levelsup), O.transform_features(d, levelsup))
        got = G.transform(d, levelsup)
        _same(got, O.transform(d, levelsup))
        if R is not None:
            _same(got, R.transform(d, levelsup))
    assert G.launch_count > 0
    G.close()


def test_gpu_equals_golden():
    g = np.load(GOLDEN)
    for c in range(int(g["ncases"])):
        k, L, seed, scoring, weighting, levelsup, n = (int(x) for x in g[f"case{c}"])
        voc, d = bow_case(k, L, seed, n=n)
        G = ORBVocabulary.from_arrays(k, L, voc["parent"], voc["descriptors"], voc["weight"], voc["word_id"], scoring, weighting)
        for name, arr in zip(("bow_word", "bow_value", "fv_node", "fv_off", "fv_feat"), G.transform(d, levelsup)):
            assert arr.tobytes() == g[f"c{c}_{name}"].tobytes(), (c, name)
        G.close()


def test_edge_cases(tmp_path):
    voc, d = bow_case(5, 4, 21, n=800, early=0.3)      # leaves above the FeatureVector level, stop words
    G = ORBVocabulary.from_arrays(5, 4, voc["parent"], voc["descriptors"], voc["weight"], voc["word_id"])
    O = orbref.RefVocabulary(arrays=voc)
    for levelsup in (0, 1, 2, 5):
        _same(G.transform_features(d, levelsup), O.transform_features(d, levelsup))
        _same(G.transform(d, levelsup), O.transform(d, levelsup))
    for n in (0, 1, 2, 31, 33, 1023, 1025):            # empty and ragged sizes around the sort's power-of-two padding
        dd = np.concatenate([d, d])[:n]
        _same(G.transform(dd, 2), O.transform(dd, 2))
    big = synth_bow_descriptors(voc, 5000, 5)          # > 4096 features: keys sorted in global scratch
    _same(G.transform(big, 2), O.transform(big, 2))
    same = np.repeat(d[:1], 700, axis=0)               # every feature on one word: one long sequential double sum
    _same(G.transform(same, 1), O.transform(same, 1))
    G.close()
    assert not ORBVocabulary().loadFromTextFile(tmp_path / "missing.txt")
    bad = tmp_path / "bad.txt"
    bad.write_text("30 3 0 0\n")
    assert not ORBVocabulary().loadFromTextFile(bad)


def test_batch_of_frames_matches_single_calls():
    voc, _ = bow_case(10, 3, 1)
    G = ORBVocabulary.from_arrays(10, 3, voc["parent"], voc["descriptors"], voc["weight"], voc["word_id"])
    O = orbref.RefVocabulary(arrays=voc)
    counts = np.array([700, 0, 1, 512, 699], np.int32)
    batch = np.zeros((len(counts), 700, 32), np.uint8)
    for b, n in enumerate(counts):
        batch[b, :n] = synth_bow_descriptors(voc, int(n), 100 + b) if n else 0
    res = G.transform_batch(batch, counts, 2)
    for b, n in enumerate(counts):
        _same(res[b], O.transform(batch[b, :n], 2))
    G.close()


def test_device_resident_batch():
    import torch
    voc, _ = bow_case(10, 3, 1)
    G = ORBVocabulary.from_arrays(10, 3, voc["parent"], voc["descriptors"], voc["weight"], voc["word_id"], device=0)
    O = orbref.RefVocabulary(arrays=voc)
    B, S = 6, 640
    counts = np.array([640, 3, 500, 0, 639, 128], np.int32)
    host = np.zeros((B, S, 32), np.uint8)
    for b, n in enumerate(counts):
        if n:
            host[b, :n] = synth_bow_descriptors(voc, int(n), 200 + b)
    dev = torch.device("cuda:0")
    d_desc, d_n = torch.from_numpy(host).to(dev), torch.from_numpy(counts).to(dev)
    bw, fn, ff = (torch.zeros((B, S), dtype=torch.int32, device=dev) for _ in range(3))
    bv = torch.zeros((B, S), dtype=torch.float64, device=dev)
    fo = torch.zeros((B, S + 1), dtype=torch.int32, device=dev)
    nb, nf = torch.zeros(B, dtype=torch.int32, device=dev), torch.zeros(B, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    G.transform_device(d_desc.data_ptr(), d_n.data_ptr(), B, S, 2, S, bw.data_ptr(), bv.data_ptr(), nb.data_ptr(), fn.data_ptr(),
                       fo.data_ptr(), ff.data_ptr(), nf.data_ptr())
    G.sync()
    for b, n in enumerate(counts):
        want = O.transform(host[b, :n], 2)
        k, g = int(nb[b]), int(nf[b])
        got = (bw[b, :k].cpu().numpy(), bv[b, :k].cpu().numpy(), fn[b, :g].cpu().numpy(), fo[b, : g + 1].cpu().numpy(),
               ff[b, : int(fo[b, g])].cpu().numpy())
        _same(got, want)
    G.close()


def test_feature_vector_feeds_search_by_bow():
    """The FeatureVector CSR is the form orb_search_by_bow consumes: merge-join of two frames' vectors (ORBmatcher.cc:180-264)."""
    voc, _ = bow_case(10, 3, 1)
    G = ORBVocabulary.from_arrays(10, 3, voc["parent"], voc["descriptors"], voc["weight"], voc["word_id"])
    a = synth_bow_descriptors(voc, 400, 1)
    fa, fb = G.transform(a, 2), G.transform(a[::-1].copy(), 2)
    common = np.intersect1d(fa[2], fb[2])
    assert len(common) > 0 and np.array_equal(fa[2], fb[2])
    G.close()
