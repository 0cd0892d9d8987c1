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
