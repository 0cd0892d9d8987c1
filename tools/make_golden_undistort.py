#!/usr/bin/env python3
"""Golden vectors for Frame::UndistortKeyPoints from cv2.undistortPoints (the only OpenCV on this box):
tests/golden/cv2_undistort.npz.  Cameras are the reference's own example settings."""
import os

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# (name, fx, fy, cx, cy, dist, width, height): ORB_SLAM2/Examples/Monocular/TUM1.yaml:8-17, TUM2.yaml, EuRoC.yaml:8-16,
# and TUM1 without k3 (a vector for which the iteration diverges near the corners: the icdist<0 exit)
CAMS = [("tum1", 517.306408, 516.469215, 318.643040, 255.313989, [0.262383, -0.953104, -0.005358, 0.002628, 1.163314], 640, 480),
        ("tum2", 520.908620, 521.007327, 325.141442, 249.701764, [0.231222, -0.784899, -0.003257, -0.000105, 0.917205], 640, 480),
        ("euroc", 458.654, 457.296, 367.215, 248.375, [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], 752, 480),
        ("tum1_nok3", 517.306408, 516.469215, 318.643040, 255.313989, [0.262383, -0.953104, -0.005358, 0.002628], 640, 480)]


def points(seed, w, h, n=3000):
    rng = np.random.default_rng(seed)
    p = np.stack([rng.uniform(0, w, n), rng.uniform(0, h, n)], 1).astype(np.float32)
    p[:4] = [[0, 0], [w, 0], [0, h], [w, h]]     # Frame::ComputeImageBounds' corners
    return p


if __name__ == "__main__":
    out = {"cv2_version": np.array(cv2.__version__)}
    for i, (name, fx, fy, cx, cy, dist, w, h) in enumerate(CAMS):
        K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
        p = points(i, w, h)
        out[name] = cv2.undistortPoints(p.reshape(-1, 1, 2), K, np.asarray(dist, np.float32), None, K).reshape(-1, 2)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "cv2_undistort.npz"), **out)
    print("wrote tests/golden/cv2_undistort.npz")
