"""GPU parity for the Hamming primitives (ref: ORB_SLAM2/src/ORBmatcher.cc:1650-1666 and the best/second scan
of :205-226): bit-exact against the oracle, plus size-independent properties at larger sizes."""
import numpy as np
import pytest

from ar_orbslam2_b200.synth import synth_descriptors, synth_queries

pytestmark = pytest.mark.gpu


def test_descriptor_distance(oracle, gpu_lib):
    from ar_orbslam2_b200.matcher import DescriptorDistance
    a = synth_descriptors(5000, 1)
    b = synth_descriptors(5000, 2)
    b[:100] = a[:100]
    b[100:200] = ~a[100:200]
    d = DescriptorDistance(a, b)
    assert np.array_equal(d, oracle.descriptor_distance(a, b))
    assert (d[:100] == 0).all() and (d[100:200] == 256).all()
    assert np.array_equal(d, DescriptorDistance(b, a))   # symmetry


@pytest.mark.parametrize("nq,ndb", [(1, 1), (3, 2), (257, 1000), (2000, 20000), (100, 33333)])
def test_top2_matches_oracle(oracle, gpu_lib, nq, ndb):
    from ar_orbslam2_b200.matcher import hamming_top2
    db = synth_descriptors(ndb, 7)
    q = synth_queries(lambda idx: db[idx], ndb, nq, 3)
    db[ndb // 2] = db[0]   # a duplicated row: ties must resolve to the first
    got = hamming_top2(q, db, idx_base=1000)
    want = oracle.hamming_top2(q, db, idx_base=1000)
    for g, w, name in zip(got, want, ("best_dist", "best_idx", "second_dist", "second_idx")):
        assert np.array_equal(g, w), name


def test_top2_empty_and_single_row(oracle, gpu_lib):
    from ar_orbslam2_b200.matcher import hamming_top2
    q = synth_descriptors(5, 1)
    d1, i1, d2, i2 = hamming_top2(q, np.zeros((0, 32), np.uint8))
    assert (d1 == 256).all() and (i1 == -1).all() and (d2 == 256).all() and (i2 == -1).all()
    d1, i1, d2, i2 = hamming_top2(q, q[:1])
    assert d1[0] == 0 and (i1 == 0).all() and (i2 == -1).all() and (d2 == 256).all()


def test_sharded_db_equals_single_scan(oracle, gpu_lib):
    """Keyframe-range shards + merge == one scan; the shards are filled on the device by the
    same generator as synth_descriptors."""
    from ar_orbslam2_b200.matcher import DescriptorDB, hamming_top2, top2_merge
    n, nq, seed = 60000, 500, 99
    full = synth_descriptors(n, seed)
    q = synth_queries(lambda idx: full[idx], n, nq, 5)
    cuts = [0, 11111, 30000, 30001, n]
    parts = []
    for a, b in zip(cuts[:-1], cuts[1:]):
        db = DescriptorDB(a, b)
        db.fill_synthetic(seed)
        assert np.array_equal(db.row(a), full[a]) and np.array_equal(db.row(b - 1), full[b - 1])
        parts.append(db.query(q))
        db.close()
    merged = top2_merge(*[np.stack([p[k] for p in parts]) for k in range(4)])
    single = hamming_top2(q, full)
    want = oracle.hamming_top2(q, full, fast=True, threads=8)
    for m, s, w in zip(merged, single, want):
        assert np.array_equal(m, s) and np.array_equal(m, w)
    # planted queries (even rows are database rows with <= 40 flipped bits) are found
    assert (merged[0][0::2] <= 40).all()


def test_top2_properties_large(gpu_lib):
    """2000 queries x 2M rows: too large for the scalar oracle; check properties instead."""
    from ar_orbslam2_b200.matcher import DescriptorDB, DescriptorDistance
    n, nq = 2_000_000, 2000
    db = DescriptorDB(0, n)
    db.fill_synthetic(5)
    q = synth_queries(lambda idx: db.rows(idx), n, nq, 11)
    d1, i1, d2, i2 = db.query(q)
    assert (d1 <= d2).all() and (i1 != i2).all() and (i1 >= 0).all() and (i2 >= 0).all()
    assert ((d1 < d2) | (i1 < i2)).all()          # ties keep scan order
    assert np.array_equal(DescriptorDistance(q, db.rows(i1)), d1)   # reported distances are real
    assert np.array_equal(DescriptorDistance(q, db.rows(i2)), d2)
    assert (d1[0::2] <= 40).all()
    db.close()


def test_popc_rate(gpu_lib):
    from ar_orbslam2_b200.matcher import measure_popc_rate
    r = measure_popc_rate()
    assert 500 < r < 50000, r


@pytest.mark.parametrize("n_kf,kf_rows,nq", [(300, 173, 500), (12, 2000, 700), (1, 5, 3), (40, 256, 257)])
def test_db_keyframe_counts(oracle, gpu_lib, n_kf, kf_rows, nq):
    """orb_db_query_keyframes: per-keyframe accepted-match counts (ORBmatcher.cc:216-231 per (query, keyframe)) and the
    shard top-2, against the oracle; keyframe boundaries fall inside the kernel's 256-row tiles."""
    from ar_orbslam2_b200.matcher import DescriptorDB
    n = n_kf * kf_rows
    full = synth_descriptors(n, 5)
    if kf_rows > 3:
        o = kf_rows if n_kf > 1 else 0
        full[o + 3] = full[o + 2]                                         # duplicate rows inside a keyframe
    q = synth_queries(lambda idx: full[idx], n, nq, 8)
    base = 7 * kf_rows                                                    # a shard that does not start at row 0
    db = DescriptorDB(base, base + n)
    db.upload(base, full)
    d1, i1, d2, i2, counts = db.query_keyframes(q, kf_rows, 50, 0.75)
    p1, pi1, p2, pi2 = db.query(q)
    assert np.array_equal(d1, p1) and np.array_equal(i1, pi1) and np.array_equal(d2, p2) and np.array_equal(i2, pi2)
    want = oracle.db_keyframe_counts(q, full, kf_rows, 50, 0.75, threads=8)
    assert np.array_equal(counts, want), np.flatnonzero(counts != want)[:8]
    assert counts.sum() >= nq // 2 - 2 or kf_rows < 10
    with pytest.raises(Exception):
        DescriptorDB(1, 1 + n).query_keyframes(q, kf_rows)               # shard not on a keyframe boundary
    db.close()
