"""OpenCV's cv::ORB with Harris score (Marker::setTargetImage / Match, ORB_SLAM2/src/Marker.cc:83-84,107-108):
the oracle restatement against cv2 4.13's golden outputs (tests/golden/cv2_orb.npz, tools/make_golden_cvorb.py) -- key
points with every field, in cv2's order, and descriptors -- and, when cv2 is importable, against cv2 live."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from make_golden_cvorb import CASES, PRIMITIVE_CASES, SWEEP, image, orb_kwargs, result_sha, sha  # noqa: E402
from ar_orbslam2_b200.synth import synth_frame  # noqa: E402
from oracle import orbref  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "cv2_orb.npz")


def oracle_orb(img, kw):
    k = orb_kwargs(kw)
    return orbref.cv_orb(img, k["nfeatures"], k["scaleFactor"], k["nlevels"], k["edgeThreshold"], k["scoreType"], k["fastThreshold"])


@pytest.mark.parametrize("name", [n for n in CASES if n != "hd_default"])
def test_oracle_equals_cv2_golden(name):
    g = np.load(GOLDEN)
    kind, w, h, seed, kw = CASES[name]
    kps, desc = oracle_orb(image(kind, w, h, seed), kw)
    assert kps.tobytes() == g[f"{name}_kps"].tobytes(), name
    assert np.array_equal(desc, g[f"{name}_desc"]), name


def test_oracle_equals_cv2_golden_hd():
    g = np.load(GOLDEN)
    kind, w, h, seed, kw = CASES["hd_default"]
    kps, desc = oracle_orb(image(kind, w, h, seed), kw)
    assert kps.tobytes() == g["hd_default_kps"].tobytes() and np.array_equal(desc, g["hd_default_desc"])


def test_primitives_equal_cv2_golden():
    g = np.load(GOLDEN)
    for w, h, seed in PRIMITIVE_CASES:
        img = synth_frame(w, h, seed)
        assert sha(orbref.resize_linear_exact(img, int(round(w / 1.2)), int(round(h / 1.2)))) == str(g[f"resize_exact_{w}x{h}_sha256"])
        assert sha(orbref.resize_linear_exact(img, w // 3 + 1, h // 2 + 3)) == str(g[f"resize_exact_odd_{w}x{h}_sha256"])
        assert sha(orbref.gaussian_blur7_sep32f(img)) == str(g[f"sep32f_blur_{w}x{h}_sha256"])


def test_oracle_equals_cv2_live():
    cv2 = pytest.importorskip("cv2")
    cv2.setNumThreads(1)
    for w, h, seed, kw in [(640, 480, 21, {}), (500, 375, 22, dict(nfeatures=800)), (752, 480, 23, dict(nlevels=4))]:
        img = synth_frame(w, h, seed)
        kps, desc = cv2.ORB_create(**orb_kwargs(kw)).detectAndCompute(img, None)
        rk, rd = oracle_orb(img, kw)
        assert len(kps) == len(rk)
        got = np.array([(p.pt[0], p.pt[1], p.size, p.angle, p.response, p.octave, p.class_id) for p in kps], dtype=rk.dtype)
        assert got.tobytes() == rk.tobytes()
        assert np.array_equal(desc, rd)


def test_oracle_equals_cv2_sweep_hashes():
    """40 further frames: SHA-256 of cv2's key points + descriptors (default parameters)"""
    g = np.load(GOLDEN)
    for (w, h, seed), want in zip(SWEEP, g["sweep_sha256"]):
        assert result_sha(*oracle_orb(synth_frame(w, h, seed), {})) == str(want), (w, h, seed)
