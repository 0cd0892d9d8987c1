"""GPU parity of orb_search_matches_ratio (Marker::searchMatches / naive_nn_search2) against the oracle."""
import numpy as np
import pytest

from ar_orbslam2_b200.matcher import searchMatches
from oracle import orbref

pytestmark = pytest.mark.gpu


def marker_case(seed, n_train=300, n_query=1200):
    rng = np.random.default_rng(seed)
    train = rng.integers(0, 256, (n_train, 32), dtype=np.uint8)
    if n_train > 10:
        train[5] = train[4]                                  # duplicate train rows: best == second -> ratio test fails
    query = rng.integers(0, 256, (n_query, 32), dtype=np.uint8)
    if n_train:
        src = rng.integers(0, n_train, n_query // 2)
        bits = np.unpackbits(train[src], axis=1)
        for i in range(len(src)):
            bits[i, rng.choice(256, int(rng.integers(0, 70)), replace=False)] ^= 1
        query[: len(src)] = np.packbits(bits, axis=1)
    return train, query


@pytest.mark.parametrize("seed,nt,nq", [(0, 300, 1200), (1, 1, 50), (2, 2, 64), (3, 0, 10), (4, 1000, 3000), (5, 257, 1)])
def test_gpu_equals_oracle(seed, nt, nq):
    train, query = marker_case(seed, nt, nq)
    qi, ti, d = searchMatches(train, query)
    oi, od = orbref.search_matches_ratio(train, query)
    keep = np.nonzero(oi >= 0)[0]
    assert np.array_equal(qi, keep) and np.array_equal(ti, oi[keep]) and np.array_equal(d, od[keep].astype(np.float32))
    if nt >= 2:
        assert len(keep) > 0 or nq < 20


@pytest.mark.parametrize("seed,nt,nq", [(10, 500, 2000), (11, 3, 100), (12, 1, 7)])
def test_gpu_equals_reference_marker_cc(seed, nt, nq):
    """liborbgpu directly against the reference's compiled Marker::searchMatches (oracle/_ref/libmarker_ref.so travels to the
    GPU box as a prebuilt file)."""
    if not orbref.ref_marker_available():
        pytest.skip("oracle/_ref/libmarker_ref.so not built")
    train, query = marker_case(seed, nt, nq)
    qi, ti, d = searchMatches(train, query)
    rq, rt, rd = orbref.ref_marker_search_matches(train, query)
    assert np.array_equal(qi, rq) and np.array_equal(ti, rt) and np.array_equal(d, rd)


@pytest.mark.parametrize("seed,nt,nq", [(20, 300, 1200), (21, 1, 50), (22, 0, 10), (23, 40, 0), (24, 1000, 33)])
def test_minmax_over_all_pairs(seed, nt, nq):
    """naive_nn_search2's minD / maxD bookkeeping (AR-1.3/src/ORBMatcher.cpp:92-93): min and max distance over every pair"""
    from ar_orbslam2_b200.matcher import hamming_minmax
    train, query = marker_case(seed, nt, max(nq, 1))
    query = query[:nq]
    got = hamming_minmax(train, query)
    if nt == 0 or nq == 0:
        assert got == (0x7fffffff, -1)
        return
    d = np.unpackbits(query[:, None, :] ^ train[None, :, :], axis=2).sum(axis=2)   # numpy restatement of hamdist2 over all pairs
    assert got == (int(d.min()), int(d.max()))
    bd, _, _, _ = orbref.hamming_top2(query, train)
    assert got[0] == int(bd.min())
