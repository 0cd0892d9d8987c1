#!/usr/bin/env python3
"""Step time of the bench's 288-frame step (extraction of 144 left + 144 right frames, then ComputeStereoMatches) for several sizes
of the extractor's L2 sub-batches (ORBGPU_CHUNK_MB; 0 = the whole batch at once), with a hash of every output so that the
schedules are seen to give the same bytes.  usage: python tools/chunk_ab.py [mb,mb,...] [width height nfeatures pairs]"""
import hashlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from ar_orbslam2_b200.extractor import ORBextractor
from ar_orbslam2_b200.stereo import stereo_match_batch_device
from ar_orbslam2_b200.synth import synth_frame

mbs = [int(a) for a in (sys.argv[1] if len(sys.argv) > 1 else "0,24,48,96").split(",")]
w, h, nf, B = (int(a) for a in sys.argv[2:6]) if len(sys.argv) >= 6 else (1241, 376, 2000, 144)
nb = 2 * B
MBF, FX = 386.1448, 718.856
frames = np.stack([synth_frame(w, h, s % 16) for s in range(B)] + [np.roll(synth_frame(w, h, s % 16), -9, axis=1) for s in range(B)])
d = torch.from_numpy(frames).cuda()
for mb in mbs:
    os.environ["ORBGPU_CHUNK_MB"] = str(mb)
    ex = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=nb)
    cap = ex.default_capacity
    kp = torch.zeros(nb * cap * 28, dtype=torch.uint8, device="cuda")
    desc = torch.zeros(nb * cap * 32, dtype=torch.uint8, device="cuda")
    n = torch.zeros(nb, dtype=torch.int32, device="cuda")
    ur = torch.zeros(B * cap, dtype=torch.float32, device="cuda")
    dp = torch.zeros(B * cap, dtype=torch.float32, device="cuda")
    nm = torch.zeros(B, dtype=torch.int32, device="cuda")
    mb_f = float(np.float32(MBF) / np.float32(FX))

    def step():
        ex.extract_batch_device(d.data_ptr(), nb, w, h, w, w * h, kp.data_ptr(), desc.data_ptr(), cap, n.data_ptr())
        stereo_match_batch_device(ex, 0, ex, B, B, kp.data_ptr(), desc.data_ptr(), n.data_ptr(), kp[B * cap * 28:].data_ptr(),
                                  desc[B * cap * 32:].data_ptr(), n[B:].data_ptr(), cap, MBF, mb_f, ur.data_ptr(), dp.data_ptr(), nm.data_ptr())

    for _ in range(3):
        step()
    ex.sync()
    st = torch.cuda.ExternalStream(ex.stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = int(os.environ.get("CHUNK_AB_REPS", "20"))
    with torch.cuda.stream(st):
        e0.record()
        for _ in range(reps):
            step()
        e1.record()
    ex.sync()
    total = e0.elapsed_time(e1) / reps
    hh = hashlib.sha256()
    nn = n.cpu().numpy()
    kpc, dc = kp.cpu().numpy().reshape(nb, cap, 28), desc.cpu().numpy().reshape(nb, cap, 32)
    for b in range(nb):
        hh.update(kpc[b, :nn[b]].tobytes()); hh.update(dc[b, :nn[b]].tobytes())
    hh.update(nn.tobytes()); hh.update(nm.cpu().numpy().tobytes())
    urc = ur.cpu().numpy().reshape(B, cap)
    for b in range(B):
        hh.update(urc[b, :nn[b]].tobytes())
    print(f"chunk_mb={mb}: {w}x{h} x{nb}: step {total:.4f} ms ({nb / total:.1f} k frames/s), keypoints {int(nn.sum())}, "
          f"launches/step {ex.launch_count // (reps + 3)}, sha256 {hh.hexdigest()[:16]}", flush=True)
    ex.close()
