#!/usr/bin/env python3
"""Stage times of one large batch (serial schedule, the library's stage events), per-cell FAST kernel vs k_fast_rows on every level vs
the default hybrid, plus the overlapped whole-step time of each.  usage: python tools/fast_ab.py [width height nfeatures batch]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from ar_orbslam2_b200 import _lib
from ar_orbslam2_b200.extractor import ORBextractor
from ar_orbslam2_b200.synth import synth_frame

w, h, nf, nb = (int(a) for a in sys.argv[1:5]) if len(sys.argv) >= 5 else (1241, 376, 2000, 288)
L = _lib.lib()
frames = np.stack([synth_frame(w, h, s % 16) for s in range(nb)])
d = torch.from_numpy(frames).cuda()
names = ["h2d", "pyramid", "fast", "octree", "blur+describe", "d2h"]
for mode in os.environ.get("FAST_AB_MODES", "cells,rows,hybrid").split(","):
    os.environ.pop("ORBGPU_FAST_ROWS_MIN_PX", None)
    os.environ.pop("ORBGPU_FAST_CELLS", None)
    if mode == "rows":
        os.environ["ORBGPU_FAST_ROWS_MIN_PX"] = "0"
    elif mode.startswith("min"):
        os.environ["ORBGPU_FAST_ROWS_MIN_PX"] = mode[3:]
    elif mode == "cells":
        os.environ["ORBGPU_FAST_CELLS"] = "1"
    ex = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=nb)
    cap = nf + 200
    kp = torch.empty(nb * cap * 28, dtype=torch.uint8, device="cuda")
    desc = torch.empty(nb * cap * 32, dtype=torch.uint8, device="cuda")
    n = torch.empty(nb, dtype=torch.int32, device="cuda")
    def step():
        ex.extract_batch_device(d.data_ptr(), nb, w, h, w, w * h, kp.data_ptr(), desc.data_ptr(), cap, n.data_ptr())
    for _ in range(3):
        step()
    ex.sync()
    st = torch.cuda.ExternalStream(ex.stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        e0.record()
        for _ in range(10):
            step()
        e1.record()
    ex.sync()
    total = e0.elapsed_time(e1) / 10
    tot_n = int(n.sum().item())
    L.orb_extractor_set_profiling(ex._h, 1)
    for _ in range(5):
        step()
    ex.sync()
    ms = (C.c_double * 6)(); calls = C.c_int64()
    L.orb_extractor_stage_times(ex._h, ms, C.byref(calls))
    per = {names[i]: round(ms[i] / max(calls.value, 1), 4) for i in range(6)}
    print(f"{mode}: {w}x{h} x{nb}: step {total:.4f} ms ({nb / total:.1f} k frames/s), keypoints {tot_n}, serial stages {per}", flush=True)
    ex.close()
