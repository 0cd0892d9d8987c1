"""CPU tests (no GPU): the oracle against the committed golden vectors.

* tests/golden/cv2_primitives.npz -- outputs of OpenCV (cv2 4.13.0) and of the host libm for the primitives whose
  source is not in /root/reference (made by tools/make_golden_cv2.py); they pin the oracle's models.
* tests/golden/oracle_hashes.json -- regression hashes of the oracle's full outputs on the benchmark frames.
"""
import hashlib
import json
import os

import numpy as np
import pytest

from ar_orbslam2_b200.synth import synth_descriptors, synth_frame, synth_stereo_pair

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "cv2_primitives.npz"))


def test_resize_matches_cv2(oracle, gold):
    cur = gold["resize_src"]
    for i in range(1, 5):
        want = gold[f"resize_l{i}"]
        cur = oracle.resize_linear(cur, want.shape[1], want.shape[0])
        assert np.array_equal(cur, want), f"cascade level {i}"
    for j in range(3):
        want = gold[f"resize_x{j}_dst"]
        assert np.array_equal(oracle.resize_linear(gold[f"resize_x{j}_src"], want.shape[1], want.shape[0]), want)


def test_border_matches_cv2(oracle, gold):
    assert np.array_equal(oracle.make_border(gold["border_src"], 19), gold["border_dst"])


def test_blur_matches_cv2(oracle, gold):
    for j in range(3):
        assert np.array_equal(oracle.gaussian_blur7(gold[f"blur{j}_src"]), gold[f"blur{j}_dst"]), j
    # the OpenCV-2.4 tap variant is selectable and differs
    assert not np.array_equal(oracle.gaussian_blur7(gold["blur0_src"], taps=1), gold["blur0_dst"])


def test_fast_matches_cv2(oracle, gold):
    for j in range(4):
        for t in (20, 7):
            got = oracle.fast(gold[f"fast{j}_src"], t)
            want = gold[f"fast{j}_t{t}"]
            assert len(got) == len(want), (j, t)
            assert np.array_equal(np.stack([got["x"], got["y"], got["response"]], 1).astype(np.int32).reshape(-1, 3), want), (j, t)


def test_fast_atan2_matches_cv2(oracle, gold):
    got = oracle.fast_atan2(gold["atan2_y"], gold["atan2_x"])
    assert np.array_equal(got.view(np.uint32), gold["atan2_deg"].view(np.uint32))
    assert oracle.fast_atan2(np.zeros(1, np.float32), np.zeros(1, np.float32))[0] == 0


def test_sincos_matches_host_libm(oracle, gold):
    s, c = oracle.sincos(gold["trig_rad"], 0)
    assert np.array_equal(s.view(np.uint32), gold["trig_sin"].view(np.uint32))
    assert np.array_equal(c.view(np.uint32), gold["trig_cos"].view(np.uint32))


def test_tables(oracle):
    """ref ORBextractor.cc:404-460; values of SURVEY 8(a) row E1."""
    t = oracle.RefExtractor(1000).tables()
    assert t["per_level"].tolist() == [217, 181, 151, 126, 105, 87, 73, 60]
    assert oracle.RefExtractor(2000).tables()["per_level"].tolist() == [434, 362, 302, 251, 209, 175, 145, 122]
    assert t["umax"].tolist() == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert t["scale"][1] == np.float32(1.2) and abs(t["scale"][7] - 3.5831816) < 1e-6
    assert np.array_equal(t["inv_scale"], (np.float32(1) / t["scale"]).astype(np.float32))


@pytest.mark.parametrize("name,w,h,nf,seed", [("C1_640x480_1000_s0", 640, 480, 1000, 0), ("C2_1241x376_2000_s3", 1241, 376, 2000, 3),
                                               ("C3_752x480_1000_s5", 752, 480, 1000, 5)])
def test_extractor_regression_hashes(oracle, name, w, h, nf, seed):
    want = json.load(open(os.path.join(HERE, "golden", "oracle_hashes.json")))["hashes"][name]
    img = synth_frame(w, h, seed)
    assert hashlib.sha256(img.tobytes()).hexdigest() == want["image_sha256"], "synthetic generator changed"
    ex = oracle.RefExtractor(nf)
    k, d = ex.extract(img)
    assert len(k) == want["n"] and [int((k["octave"] == l).sum()) for l in range(8)] == want["per_level"]
    assert [len(ex.candidates(l)) for l in range(8)] == want["candidates"]
    assert hashlib.sha256(k.tobytes()).hexdigest() == want["keypoints_sha256"]
    assert hashlib.sha256(d.tobytes()).hexdigest() == want["descriptors_sha256"]
    assert hashlib.sha256(b"".join(ex.pyramid_level(l, True).tobytes() for l in range(8))).hexdigest() == want["pyramid_sha256"]
    # structural properties of the reference's output
    assert (np.diff(k["octave"]) >= 0).all(), "keypoints are level-major"
    assert (k["class_id"] == -1).all() and (k["angle"] >= 0).all() and (k["angle"] < 360).all()
    sizes = (31 * ex.tables()["scale"]).astype(np.int32).astype(np.float32)
    assert np.array_equal(k["size"], sizes[k["octave"]])
    # libm mode (the reference's literal cosf/sinf calls) gives the same descriptors
    ex.set_trig_mode(1)
    k1, d1 = ex.extract(img)
    assert np.array_equal(d, d1) and k.tobytes() == k1.tobytes()


def test_stereo_regression_hash(oracle):
    want = json.load(open(os.path.join(HERE, "golden", "oracle_hashes.json")))["hashes"]["C2_stereo_s0"]
    l, r = synth_stereo_pair(1241, 376, 0)
    a, b = oracle.RefExtractor(2000), oracle.RefExtractor(2000)
    kl, dl = a.extract(l)
    kr, dr = b.extract(r)
    rc, ur, dp, sad, nm = oracle.stereo_match(a, b, kl, dl, kr, dr, 386.1448, np.float32(386.1448) / np.float32(718.856))
    assert rc == want["rc"] and nm == want["n_matched"]
    assert hashlib.sha256(ur.tobytes()).hexdigest() == want["uright_sha256"]
    assert hashlib.sha256(dp.tobytes()).hexdigest() == want["depth_sha256"]
    m = ur != -1
    assert (dp[m] > 0).all() and np.allclose(dp[m], np.float32(386.1448) / (kl["x"][m] - ur[m]), rtol=1e-6)


def test_extractor_edge_cases(oracle):
    ex = oracle.RefExtractor(500)
    rc, k, d = ex.extract(np.zeros((0, 0), np.uint8), check=False)
    assert rc == 0 and len(k) == 0
    rc, k, d = ex.extract(synth_frame(160, 120, 1), check=False)
    assert rc == -3, "level 7 of a 160x120 image is under the 30-px cell grid: the reference divides by zero"
    k, d = ex.extract(np.full((240, 320), 9, np.uint8))
    assert len(k) == 0
    rc, k, d = ex.extract(synth_frame(640, 480, 2), capacity=10, check=False)
    assert rc == -4


def test_hamming_oracle_properties(oracle):
    a, b = synth_descriptors(300, 1), synth_descriptors(300, 2)
    d = oracle.descriptor_distance(a, b)
    bits = np.unpackbits(a ^ b, axis=1).sum(1)
    assert np.array_equal(d, bits)
    d1, i1, d2, i2 = oracle.hamming_top2(a[:50], b)
    full = np.unpackbits(a[:50, None, :] ^ b[None, :, :], axis=2).sum(2)
    assert np.array_equal(d1, full.min(1)) and np.array_equal(i1, full.argmin(1))
    f1, j1, f2, j2 = oracle.hamming_top2(a[:50], b, fast=True, threads=4)
    assert np.array_equal(d1, f1) and np.array_equal(i1, j1) and np.array_equal(d2, f2) and np.array_equal(i2, j2)


def test_rotation_histogram_quirk(oracle):
    """factor = 1/HISTO_LENGTH (ORBmatcher.cc:172): 30-degree bins, only bins 0..12 reachable."""
    import ctypes as C
    L = oracle.lib()
    bins = {L.orbref_rot_bin(C.c_float(a), C.c_float(0.0)) for a in np.arange(0, 360, 0.25)}
    assert bins == set(range(13))
    assert L.orbref_rot_bin(C.c_float(10.0), C.c_float(350.0)) == 1   # -340 + 360 = 20 deg -> round(0.667) = 1


def test_sweep_hashes_reproduce(oracle):
    """tests/golden/oracle_sweep_hashes.json (the GPU box's yardstick for the seed sweeps) is what the oracle produces here:
    a sample of every section is regenerated."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden_sweeps", os.path.join(HERE, "..", "tools", "make_golden_sweeps.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    want = json.load(open(os.path.join(HERE, "golden", "oracle_sweep_hashes.json")))
    assert len(want["C1"]) == 100 and len(want["C2"]) == 32 and len(want["C4"]) == 16 and len(want["bench"]) == 8
    for s in (0, 41, 99):
        assert mod.one_frame((640, 480, 1000, s)) == want["C1"][s]
    assert mod.one_pair((1241, 376, 2000, 13)) == want["C2"][13]
    assert mod.one_frame((1920, 1080, 2000, 7)) == want["C4"][7]
    assert mod.one_pair((1241, 376, 2000, 17 + 3000 + 5)) == want["bench"]["3"][5]
