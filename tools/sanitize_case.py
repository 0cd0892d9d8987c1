#!/usr/bin/env python3
"""Smallest end-to-end case for compute-sanitizer: one stereo frame through extraction, stereo association, the tracking
matchers and a tiny brute-force scan.  usage: compute-sanitizer --tool memcheck python tools/sanitize_case.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ar_orbslam2_b200.extractor import ORBextractor
from ar_orbslam2_b200.matcher import hamming_top2
from ar_orbslam2_b200.stereo import ComputeStereoMatches
from ar_orbslam2_b200.synth import synth_descriptors, synth_stereo_pair
from ar_orbslam2_b200.tracking import Frame, ORBmatcher

w, h, nf = 641, 377, 600          # odd sizes on purpose
l, r = synth_stereo_pair(w, h, 3)
a = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
b = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
kl, dl = a(l); kr, dr = b(r)
ur, dp = ComputeStereoMatches(a, b, kl, dl, kr, dr, 386.1448, 0.537, allow_empty=True)
sc = a.GetScaleFactors()
f1 = Frame(kl, dl, sc, (0, w, 0, h)); f2 = Frame(kr, dr, sc, (0, w, 0, h))
prev = np.stack([kl["x"], kl["y"]], 1).astype(np.float32).copy()
n, m = ORBmatcher(0.9, True).SearchForInitialization(f1, f2, prev, 100)
n2, asg = ORBmatcher(0.8).SearchByProjectionPoints(f2, np.ones(len(kl), np.uint8), kl["x"], kl["y"], kl["x"] - 3, kl["octave"], np.full(len(kl), 0.9, np.float32), dl, 3.0)
d = hamming_top2(synth_descriptors(33, 1), synth_descriptors(1000, 2))
print("ok", len(kl), len(kr), int((ur != -1).sum()), n, n2, int(d[0].min()))
# repeated calls (graph capture + replay, per-level quadtree streams), the single-call stereo front end, a dense-noise frame
from ar_orbslam2_b200.stereo import StereoFrame
for _ in range(3):
    a(l)
c = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=2)
for _ in range(3):
    out = StereoFrame(c, l, r, 386.1448, 0.537, allow_empty=True)
noise = np.random.default_rng(1).integers(0, 256, (h, w), dtype=np.uint8)
kn, dn = a(noise)
ramp = ((np.arange(w)[None, :] * 9 + np.arange(h)[:, None] * 11) % 256).astype(np.uint8)   # steep ramps: both-type compass survivors
kq, dq = a(ramp)
print("ok2", a.graph_replays, len(out[0]), len(kn), len(kq))
# round 2: frame index built on the device, SearchByProjection(Cur, Last), BoW groups, keyframe-count scan, k_fast_rows batch
from ar_orbslam2_b200.matcher import DescriptorDB
a(l)
fd = Frame.from_extractor(a, 0, (0, w, 0, h))
n0, m0 = ORBmatcher(0.9, True).SearchForInitialization(fd, f2, prev.copy(), 100)
N = len(kl)
n3, asg3 = ORBmatcher(0.9, True).SearchByProjectionLast(f2, np.ones(N, np.uint8), kl["x"] - 2, kl["y"] + 1, np.full(N, 0.2, np.float32),
                                                        kl["octave"], kl["angle"], dl, 15.0, obs_positive=np.ones(N, np.uint8))
g = 8   # 8 vocabulary groups: feature i of either frame falls in group i % 8
ko = np.cumsum([0] + [len(range(k, len(kl), g)) for k in range(g)]).astype(np.int32)
ki = np.concatenate([np.arange(k, len(kl), g) for k in range(g)]).astype(np.int32)
fo = np.cumsum([0] + [len(range(k, len(kr), g)) for k in range(g)]).astype(np.int32)
fi = np.concatenate([np.arange(k, len(kr), g) for k in range(g)]).astype(np.int32)
n4, m4 = ORBmatcher(0.7, True).SearchByBoW(f1, f2, ko, ki, fo, fi, np.ones(len(kl), np.uint8))
db = DescriptorDB(600, 600 + 4 * 300)
db.upload(600, synth_descriptors(1200, 4))
kfres = db.query_keyframes(synth_descriptors(70, 5), 300)
db.close()
os.environ["ORBGPU_FAST_ROWS_MIN_PX"] = "0"
big = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=18)
outs = big.extract_batch(np.stack([l, r, noise, ramp] * 4 + [l, r]))
print("ok3", n0, n3, n4, int(kfres[4].sum()), sum(len(o[0]) for o in outs))
