#!/usr/bin/env python3
"""One keyframe-mode brute-force scan for an ncu capture: k_hamming_top2<true> over 2 M database rows (1000 keyframes x 2000
descriptors, 64 MB) for 2000 queries -- config C5's kernel on a slice that ncu can replay in seconds.
usage (on the GPU box): ncu --set full --clock-control none --import-source on -k regex:k_hamming_top2 -c 2 -o gpurun_out/r3_hamming
                            python tools/hamming_ncu_case.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ar_orbslam2_b200.matcher import DescriptorDB  # noqa: E402
from ar_orbslam2_b200.synth import synth_descriptors  # noqa: E402

KF, ROWS = int(os.environ.get("KEYFRAMES", "1000")), 2000
db = DescriptorDB(0, KF * ROWS, device=0)
db.fill_synthetic(7)
q = synth_descriptors(2000, 99)
q[:100] = db.rows(np.arange(100) * 1777)          # a few planted queries so the accept path is taken
for _ in range(3):
    out = db.query_keyframes(q, ROWS)
print("best distance of the planted queries:", out[0][:5], "rows", out[1][:5])
db.close()
