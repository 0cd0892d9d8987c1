"""CPU test (no GPU): the restated oracle against oracle/_ref -- the reference's OWN ORBextractor.cc compiled unmodified
against the stub OpenCV of oracle/cvstub (primitives = the cv2-pinned models, monotonic allocator so the reference's
pointer tie-break is deterministic).  Keypoints and descriptors must be byte-identical.  Skipped only where neither
the prebuilt oracle/_ref/liborbslam_ref.so nor /root/reference exists."""
import numpy as np
import pytest

from ar_orbslam2_b200.synth import synth_frame

CASES = [(640, 480, 1000, 0), (1241, 376, 2000, 3), (752, 480, 2000, 5), (640, 480, 300, 7), (320, 240, 100, 9), (400, 300, 1500, 4)]


@pytest.fixture(scope="module")
def ref(oracle):
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built and /root/reference absent")
    return oracle


@pytest.mark.parametrize("w,h,nf,seed", CASES)
def test_oracle_equals_reference_source(ref, w, h, nf, seed):
    img = synth_frame(w, h, seed)
    ka, da = ref.RefExtractor(nf).extract(img)
    r = ref.ReferenceExtractor(nf)
    kb, db = r.extract(img)
    assert len(ka) == len(kb)
    assert ka.tobytes() == kb.tobytes(), "keypoints differ from the reference source"
    assert np.array_equal(da, db), "descriptors differ from the reference source"


def test_oracle_equals_reference_source_on_noise(ref):
    rng = np.random.default_rng(1)
    for img in (rng.integers(0, 256, (480, 640), dtype=np.uint8), (128 + 25 * rng.standard_normal((376, 1241))).clip(0, 255).astype(np.uint8)):
        ka, da = ref.RefExtractor(1000).extract(img)
        kb, db = ref.ReferenceExtractor(1000).extract(img)
        assert ka.tobytes() == kb.tobytes() and np.array_equal(da, db)


def test_reference_pyramid_equals_oracle(ref):
    img = synth_frame(640, 480, 12)
    a = ref.RefExtractor(500)
    a.extract(img)
    r = ref.ReferenceExtractor(500)
    r.extract(img)
    for l in range(8):
        assert np.array_equal(a.pyramid_level(l), r.pyramid_level(l))


def _marker_case(seed, n_train, n_query, max_flips=70):
    rng = np.random.default_rng(seed)
    train = rng.integers(0, 256, (n_train, 32), dtype=np.uint8)
    if n_train > 10:
        train[5] = train[4]                       # duplicate train rows: best == second, the 0.8 ratio test must fail
        train[9] = train[8]; train[9, 0] ^= 1     # and a near-duplicate one bit away
    query = rng.integers(0, 256, (n_query, 32), dtype=np.uint8)
    if n_train:
        src = rng.integers(0, n_train, n_query // 2)
        bits = np.unpackbits(train[src], axis=1)
        for i in range(len(src)):
            bits[i, rng.choice(256, int(rng.integers(0, max_flips)), replace=False)] ^= 1
        query[: len(src)] = np.packbits(bits, axis=1)
    return train, query


@pytest.mark.parametrize("seed,nt,nq", [(0, 300, 1200), (1, 1, 50), (2, 2, 64), (3, 0, 10), (4, 1000, 2000), (5, 257, 1), (6, 64, 300)])
def test_search_matches_ratio_equals_marker_cc(oracle, seed, nt, nq):
    """orbref_search_matches_ratio (the restatement the GPU test compares with) against the reference's own
    Marker::searchMatches + hamdist2 (src/Marker.cc:314-358, compiled from the reference's file by Makefile.ref)."""
    if not oracle.ref_marker_available():
        pytest.skip("oracle/_ref/libmarker_ref.so not built and /root/reference absent")
    train, query = _marker_case(seed, nt, nq, 70 if seed != 6 else 20)
    qi, ti, d = oracle.ref_marker_search_matches(train, query)
    oi, od = oracle.search_matches_ratio(train, query)
    keep = np.nonzero(oi >= 0)[0]
    assert np.array_equal(qi, keep) and np.array_equal(ti, oi[keep]) and np.array_equal(d, od[keep].astype(np.float32))
    if nt >= 2 and nq >= 50:
        assert len(keep) > 0
