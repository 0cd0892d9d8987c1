"""OpenCV's cv::ORB (Harris score) on the GPU (Marker::setTargetImage / Match, ORB_SLAM2/src/Marker.cc:83-84,
107-108) through the C ABI: bit-identical to cv2 4.13's golden outputs (tests/golden/cv2_orb.npz) -- every cv::KeyPoint
field, cv2's key-point ORDER, every descriptor -- and to the oracle on further seeds; the parallel replay of
std::nth_element + std::partition (KeyPointsFilter::retainBest) against libstdc++ itself."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from make_golden_cvorb import CASES, SWEEP, image, orb_kwargs, result_sha  # noqa: E402
from ar_orbslam2_b200 import cvorb  # noqa: E402
from ar_orbslam2_b200.synth import synth_frame  # noqa: E402
from oracle import orbref  # noqa: E402

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "cv2_orb.npz")


def gpu_orb(kw, w, h):
    k = orb_kwargs(kw)
    return cvorb.ORB(k["nfeatures"], k["scaleFactor"], k["nlevels"], k["edgeThreshold"], 0, 2, k["scoreType"], 31, k["fastThreshold"],
                     max_width=w, max_height=h)


def oracle_orb(img, kw):
    k = orb_kwargs(kw)
    return orbref.cv_orb(img, k["nfeatures"], k["scaleFactor"], k["nlevels"], k["edgeThreshold"], k["scoreType"], k["fastThreshold"])


@pytest.mark.parametrize("name", list(CASES))
def test_gpu_equals_cv2_golden(name):
    g = np.load(GOLDEN)
    kind, w, h, seed, kw = CASES[name]
    orb = gpu_orb(kw, w, h)
    kps, desc = orb(image(kind, w, h, seed))
    assert len(kps) == len(g[f"{name}_kps"]), name
    assert kps.tobytes() == g[f"{name}_kps"].tobytes(), name
    assert np.array_equal(desc, g[f"{name}_desc"]), name
    orb.close()


def test_gpu_equals_oracle_seeds_and_sizes():
    """one handle, changing image sizes (geometry rebuilt), against the oracle; blurred layers too"""
    orb = gpu_orb({}, 1300, 800)
    for w, h, seed in [(640, 480, 40), (1241, 376, 41), (752, 480, 42), (333, 777, 43), (1300, 800, 44), (640, 480, 45), (100, 90, 46)]:
        img = synth_frame(w, h, seed)
        kps, desc = orb(img)
        rk, rd = oracle_orb(img, {})
        assert kps.tobytes() == rk.tobytes(), (w, h, seed)
        assert np.array_equal(desc, rd), (w, h, seed)
        for lvl in (0, 3, 7):
            assert np.array_equal(orb.blurred_level(lvl), orbref.cv_orb_level(img, lvl, blurred=True)), (w, h, lvl)
    # pitched view and an empty image
    big = synth_frame(800, 600, 47)
    view = big[50:530, 60:700]
    kps, desc = orb(view)
    rk, rd = oracle_orb(np.ascontiguousarray(view), {})
    assert kps.tobytes() == rk.tobytes() and np.array_equal(desc, rd)
    kps, desc = orb(np.zeros((0, 0), np.uint8))
    assert len(kps) == 0 and desc.shape == (0, 32)
    assert orb.launch_count > 0
    orb.close()


def _retain_cases():
    rng = np.random.default_rng(7)
    out = []
    for n in (1, 2, 3, 4, 5, 7, 31, 32, 33, 100, 1000, 1025, 4097, 20000, 70001):
        for kind in ("ties", "few", "distinct", "sorted", "reversed", "equal"):
            if kind == "ties":
                r = rng.integers(20, 60, n).astype(np.float32)          # FAST scores: heavy ties
            elif kind == "few":
                r = rng.integers(0, 3, n).astype(np.float32)
            elif kind == "distinct":
                r = rng.standard_normal(n).astype(np.float32) * 1e-5      # Harris responses, both signs
            elif kind == "sorted":
                r = np.arange(n, dtype=np.float32)
            elif kind == "reversed":
                r = np.arange(n, 0, -1).astype(np.float32)
            else:
                r = np.full(n, 5, np.float32)
            for npts in sorted({0, 1, 2, n // 7, n // 2, n - 1, n, n + 5}):
                if npts >= 0:
                    out.append((r, npts))
    return out


def test_retain_best_replay_equals_libstdcxx():
    """the parallel Hoare replay leaves exactly the permutation std::nth_element + std::partition leave"""
    for r, npts in _retain_cases():
        got = cvorb.retain_best(r, npts)
        want = orbref.retain_best(r, npts)
        assert np.array_equal(got, want), (len(r), npts, r[:8])


def test_retain_best_depth_limit_fallback():
    """forced depth limits drive __introselect into its heap-select fallback (libstdc++'s own __introselect is the checker)"""
    rng = np.random.default_rng(8)
    for n in (5, 17, 100, 1000, 5000):
        for kind in range(3):
            r = (rng.integers(0, 40, n) if kind == 0 else rng.standard_normal(n) if kind == 1 else np.arange(n)).astype(np.float32)
            for npts in (1, 2, n // 3, n // 2, n - 2):
                if npts < 1:
                    continue
                for depth in (0, 1, 2, 4):
                    got = cvorb.retain_best(r, npts, depth)
                    want = orbref.retain_best(r, npts, depth)
                    assert np.array_equal(got, want), (n, kind, npts, depth)


def test_gpu_equals_cv2_sweep_hashes():
    """40 further frames on one handle: SHA-256 of key points + descriptors equal cv2's"""
    g = np.load(GOLDEN)
    orb = gpu_orb({}, 752, 480)
    for (w, h, seed), want in zip(SWEEP, g["sweep_sha256"]):
        assert result_sha(*orb(synth_frame(w, h, seed))) == str(want), (w, h, seed)
    orb.close()


def test_device_resident_image_and_large_host_image():
    """image_loc = ORB_DEVICE (pitched device buffer) and a host image above the pinned-staging threshold give the host path's bytes"""
    import torch
    orb = gpu_orb({}, 1920, 1080)
    for w, h, seed in [(640, 480, 60), (1920, 1080, 61)]:
        img = synth_frame(w, h, seed)
        hk, hd = orb(img)
        pitch = (w + 255) // 256 * 256
        d = torch.zeros((h, pitch), dtype=torch.uint8, device="cuda")
        d[:, :w] = torch.from_numpy(img).cuda()
        torch.cuda.synchronize()
        dk, dd = orb.detect_and_compute_device(d.data_ptr(), w, h, pitch)
        assert dk.tobytes() == hk.tobytes() and np.array_equal(dd, hd), (w, h)
        rk, rd = oracle_orb(img, {})
        assert hk.tobytes() == rk.tobytes() and np.array_equal(hd, rd), (w, h)
    orb.close()
