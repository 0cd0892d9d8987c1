#!/usr/bin/env python3
"""Verbose stage-by-stage diff of liborbgpu vs the oracle on one frame (debug aid, run on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import orbref
from ar_orbslam2_b200.extractor import ORBextractor
from ar_orbslam2_b200.synth import synth_frame

w, h, nf, seed = (int(a) for a in (sys.argv[1:5] if len(sys.argv) >= 5 else (640, 480, 1000, 0)))
img = synth_frame(w, h, seed)
ref = orbref.RefExtractor(nf)
rk, rd = ref.extract(img)
ex = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
gk, gd = ex(img)
print("n gpu", len(gk), "n ref", len(rk))
for l in range(8):
    a, b = ex.pyramid_level(l, True), ref.pyramid_level(l, True)
    ca, cb = ex.candidates(l), ref.candidates(l)
    same = len(ca) == len(cb) and np.array_equal(ca, cb)
    print(f"L{l} size {ex.level_size(l)} pyr diff {int((a != b).sum())} cands {len(ca)}/{len(cb)} same={same} "
          f"blur diff {int((ex.blurred_level(l) != ref.blurred_level(l)).sum())} kept {int((gk['octave']==l).sum())}/{int((rk['octave']==l).sum())}")
    if not same:
        sa = set(map(tuple, ca.tolist())); sb = set(map(tuple, cb.tolist()))
        print("   only gpu", sorted(sa - sb)[:10], "only ref", sorted(sb - sa)[:10])
n = min(len(gk), len(rk))
for f in gk.dtype.names:
    bad = np.flatnonzero(gk[f][:n].view(np.uint32) != rk[f][:n].view(np.uint32))
    print(f, "mismatches", len(bad), bad[:10], gk[f][bad[:5]], rk[f][bad[:5]])
print("desc mismatching rows", int((gd[:n] != rd[:n]).any(axis=1).sum()))
