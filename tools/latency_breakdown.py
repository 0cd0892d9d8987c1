#!/usr/bin/env python3
"""Where one blocking single-frame call spends its time: wall clock of the host-array call, wall clock of the same call with the
frame and the results resident on the device (+ one synchronisation), and the device span of its launches (CUDA events on the
handle's stream).  usage: python tools/latency_breakdown.py [width height nfeatures]"""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from ar_orbslam2_b200.extractor import ORBextractor
from ar_orbslam2_b200.synth import synth_frame

w, h, nf = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (640, 480, 1000)
ex = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
img = synth_frame(w, h, 5)
cap = ex.default_capacity
d = torch.from_numpy(img).cuda()
kp = torch.zeros(cap * 28, dtype=torch.uint8, device="cuda"); desc = torch.zeros(cap * 32, dtype=torch.uint8, device="cuda")
n = torch.zeros(1, dtype=torch.int32, device="cuda")
st = torch.cuda.ExternalStream(ex.stream)


def med(f, reps=200):
    for _ in range(10):
        f()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter(); f(); t.append(time.perf_counter() - t0)
    return 1e6 * float(np.median(t))


host = med(lambda: ex(img))


def dev():
    ex.extract_batch_device(d.data_ptr(), 1, w, h, w, w * h, kp.data_ptr(), desc.data_ptr(), cap, n.data_ptr())
    ex.sync()


devw = med(dev)
spans = []
for _ in range(100):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        e0.record()
        ex.extract_batch_device(d.data_ptr(), 1, w, h, w, w * h, kp.data_ptr(), desc.data_ptr(), cap, n.data_ptr())
        e1.record()
    ex.sync()
    spans.append(1e3 * e0.elapsed_time(e1))
print(f"{w}x{h} nf={nf}: host-array blocking call {host:.1f} us | device-resident call + sync {devw:.1f} us | device span of the launches "
      f"{np.median(spans):.1f} us | graph replays {ex.graph_replays}, keypoints {int(n.item())}")
ex.close()
