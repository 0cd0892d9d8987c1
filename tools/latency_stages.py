#!/usr/bin/env python3
"""Per-stage device time of ONE frame through the blocking host API (serial schedule, stage events of the library):
where a tracking thread's per-frame wait goes.  usage: python tools/latency_stages.py [width height nfeatures]"""
import ctypes as C
import sys
import time

import numpy as np

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from ar_orbslam2_b200 import _lib
from ar_orbslam2_b200.extractor import ORBextractor
from ar_orbslam2_b200.synth import synth_frame

w, h, nf = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (640, 480, 1000)
L = _lib.lib()
ex = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
img = synth_frame(w, h, 5)
for _ in range(5):
    ex(img)
t = []
for _ in range(50):
    t0 = time.perf_counter(); ex(img); t.append(time.perf_counter() - t0)
print(f"{w}x{h} nf={nf}: blocking call median {1e3 * np.median(t):.4f} ms (graph replays {ex.graph_replays})")
L.orb_extractor_set_profiling(ex._h, 1)
for _ in range(50):
    ex(img)
ms = (C.c_double * 6)(); calls = C.c_int64()
L.orb_extractor_stage_times(ex._h, ms, C.byref(calls))
names = ["h2d", "pyramid", "fast_cells", "octree", "blur+describe", "d2h"]
print("  serial-schedule stage times, us per call:", {n: round(1e3 * ms[i] / calls.value, 1) for i, n in enumerate(names)},
      "sum", round(1e3 * sum(ms) / calls.value, 1))
L.orb_extractor_set_profiling(ex._h, 0)
ex.close()
