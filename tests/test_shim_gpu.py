"""GPU test of the C++ drop-in shim (ar_orbslam2_b200/shim/ORBextractor.{h,cc}): the reference's class interface --
ORBextractor(int,float,int,int,int), operator()(InputArray, InputArray, vector<KeyPoint>&, OutputArray), the getters,
mvImagePyramid -- driven as Frame's stereo constructor drives it (two extractors on two threads, ref Frame.cc:83-95),
compared with the oracle."""
import os
import subprocess

import numpy as np
import pytest

from ar_orbslam2_b200._abi import KEYPOINT_DTYPE
from ar_orbslam2_b200.synth import synth_stereo_pair

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_shim_stereo_frame(oracle, gpu_lib, tmp_path):
    exe = os.path.join(ROOT, "ar_orbslam2_b200", "shim", "shim_selftest")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-s", "-C", os.path.dirname(exe)])
    w, h, nf = 1241, 376, 2000
    left, right = synth_stereo_pair(w, h, 77)
    left.tofile(tmp_path / "l.raw"); right.tofile(tmp_path / "r.raw")
    out = subprocess.run([exe, str(w), str(h), str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(nf), str(tmp_path / "o")],
                         capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    assert "shim ok=1" in out.stdout and "levels=8" in out.stdout
    rl, rr = oracle.RefExtractor(nf), oracle.RefExtractor(nf)
    kl, dl = rl.extract(left)
    kr, dr = rr.extract(right)
    mbf = np.float32(386.1448)
    rc, ur, dp, _, _ = oracle.stereo_match(rl, rr, kl, dl, kr, dr, mbf, mbf / np.float32(718.856))
    g = lambda name, dt: np.fromfile(tmp_path / f"o_{name}.bin", dt)
    assert g("kl", KEYPOINT_DTYPE).tobytes() == kl.tobytes() and g("kr", KEYPOINT_DTYPE).tobytes() == kr.tobytes()
    assert np.array_equal(g("dl", np.uint8).reshape(-1, 32), dl) and np.array_equal(g("dr", np.uint8).reshape(-1, 32), dr)
    assert np.array_equal(g("ur", np.uint32), ur.view(np.uint32)) and np.array_equal(g("dp", np.uint32), dp.view(np.uint32))
    # mvImagePyramid mirror keeps the reference's padded-plane layout (19-px reflect border around the ROI)
    assert np.array_equal(g("l3", np.uint8), rl.pyramid_level(3, True).ravel())
