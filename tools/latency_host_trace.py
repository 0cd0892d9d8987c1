#!/usr/bin/env python3
"""Host-side split of one blocking single-frame call: the ctypes call itself (orb_extract into preallocated arrays) against the
Python wrapper around it.  usage: python tools/latency_host_trace.py [width height nfeatures]"""
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
cap = ex.default_capacity
kp = np.zeros(cap * 28, np.uint8); desc = np.zeros(cap * 32, np.uint8); n = C.c_int(0)
f = L.orb_extract
args = (ex._h, img.ctypes.data_as(C.c_void_p), w, h, w, 0, kp.ctypes.data_as(C.c_void_p), desc.ctypes.data_as(C.c_void_p), cap, C.byref(n))


def med(fn, reps=300):
    for _ in range(10):
        fn()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); t.append(time.perf_counter() - t0)
    return 1e6 * float(np.median(t))


print(f"{w}x{h} nf={nf}: wrapper call {med(lambda: ex(img)):.1f} us | bare orb_extract through ctypes {med(lambda: f(*args)):.1f} us (n={n.value})")
ex.close()
