#!/usr/bin/env python3
"""Generate tests/golden/cv2_primitives.npz and tests/golden/oracle_hashes.json.

The reference's arithmetic for resize / copyMakeBorder / GaussianBlur / FAST / fastAtan2 lives in OpenCV, which
is not vendored in /root/reference.  These vectors are produced by the only OpenCV on the build box
(Python cv2, version recorded in the file) from small synthetic inputs, and pin the oracle's primitive models.
The libm vectors pin the restated cosf/sinf against the host glibc.
Run from the repo root:  python tools/make_golden_cv2.py
"""
import ctypes
import hashlib
import json
import os
import sys

import cv2
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ar_orbslam2_b200.synth import synth_frame, synth_stereo_pair  # noqa: E402
from oracle import orbref  # noqa: E402

out = {}
meta = {"cv2_version": cv2.__version__, "numpy": np.__version__}

# --- resize cascade at the reference's level ratios (ORBextractor.cc:1049-1061), small image
img = synth_frame(230, 170, 901)
out["resize_src"] = img
cur = img
sizes = [(192, 142), (160, 118), (133, 98), (111, 82)]
for i, (w, h) in enumerate(sizes):
    cur = cv2.resize(cur, (w, h), interpolation=cv2.INTER_LINEAR)
    out[f"resize_l{i + 1}"] = cur
# odd aspect / near-identity / 2:1 cases
for j, (sw, sh, dw, dh) in enumerate([(61, 47, 60, 46), (64, 48, 32, 24), (50, 40, 29, 23)]):
    s = synth_frame(sw, sh, 950 + j)
    out[f"resize_x{j}_src"] = s
    out[f"resize_x{j}_dst"] = cv2.resize(s, (dw, dh), interpolation=cv2.INTER_LINEAR)

# --- copyMakeBorder 19 px REFLECT_101 (ORBextractor.cc:1063-1068)
out["border_src"] = synth_frame(71, 53, 902)
out["border_dst"] = cv2.copyMakeBorder(out["border_src"], 19, 19, 19, 19, cv2.BORDER_REFLECT_101)

# --- GaussianBlur 7x7 sigma 2 REFLECT_101 (ORBextractor.cc:1024-1026)
for j, (w, h) in enumerate([(123, 91), (40, 33)]):
    s = synth_frame(w, h, 903 + j)
    out[f"blur{j}_src"] = s
    out[f"blur{j}_dst"] = cv2.GaussianBlur(s, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
rng = np.random.default_rng(7)
s = rng.integers(0, 256, (64, 80), dtype=np.uint8)
out["blur2_src"] = s
out["blur2_dst"] = cv2.GaussianBlur(s, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)

# --- FAST-9/16 with NMS at 20 and 7 (ORBextractor.cc:776-783): texture, noise, tiny ROI
fast_imgs = [synth_frame(150, 110, 905), rng.integers(0, 256, (60, 70), dtype=np.uint8), synth_frame(46, 41, 906),
             synth_frame(9, 8, 907)]
for j, s in enumerate(fast_imgs):
    out[f"fast{j}_src"] = s
    for t in (20, 7):
        det = cv2.FastFeatureDetector_create(t, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        kps = det.detect(np.ascontiguousarray(s))
        arr = np.array([(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in kps], np.int32).reshape(-1, 3)
        for k in kps:
            assert k.size == 7 and k.angle == -1 and k.octave == 0 and k.class_id == -1
        out[f"fast{j}_t{t}"] = arr

# --- fastAtan2 (ORBextractor.cc:97) on integer-valued moments incl. axes and zero
y = rng.integers(-60000, 60000, 4000).astype(np.float32)
x = rng.integers(-60000, 60000, 4000).astype(np.float32)
y[:50] = 0; x[50:100] = 0; y[100:110] = 0; x[100:110] = 0
x[110:130] = y[110:130]; x[130:150] = -y[130:150]
out["atan2_y"], out["atan2_x"] = y, x
out["atan2_deg"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)

# --- host libm cosf/sinf on angle*pi/180 (ORBextractor.cc:100-104)
libm = ctypes.CDLL("libm.so.6")
libm.cosf.restype = ctypes.c_float; libm.cosf.argtypes = [ctypes.c_float]
libm.sinf.restype = ctypes.c_float; libm.sinf.argtypes = [ctypes.c_float]
ang = np.concatenate([rng.uniform(0, 360, 6000), np.arange(0, 360, 0.5), [0.0, 1e-3, 42.97, 89.99999, 359.99997]]).astype(np.float32)
rad = ang * np.float32(np.float64(np.pi) / np.float64(np.float32(180.0)))
out["trig_rad"] = rad.astype(np.float32)
out["trig_sin"] = np.array([libm.sinf(float(r)) for r in rad], np.float32)
out["trig_cos"] = np.array([libm.cosf(float(r)) for r in rad], np.float32)
import platform
meta["libc"] = " ".join(platform.libc_ver())

os.makedirs("tests/golden", exist_ok=True)
np.savez_compressed("tests/golden/cv2_primitives.npz", **out)

# --- regression hashes of the oracle's own outputs on the benchmark frames
hashes = {}
for name, (w, h, nf, seed) in {"C1_640x480_1000_s0": (640, 480, 1000, 0), "C2_1241x376_2000_s3": (1241, 376, 2000, 3),
                               "C3_752x480_1000_s5": (752, 480, 1000, 5), "HD_1920x1080_2000_s2": (1920, 1080, 2000, 2)}.items():
    img = synth_frame(w, h, seed)
    ex = orbref.RefExtractor(nf)
    k, d = ex.extract(img)
    hashes[name] = {"image_sha256": hashlib.sha256(img.tobytes()).hexdigest(), "n": int(len(k)),
                    "per_level": [int((k["octave"] == l).sum()) for l in range(8)],
                    "candidates": [int(len(ex.candidates(l))) for l in range(8)],
                    "keypoints_sha256": hashlib.sha256(k.tobytes()).hexdigest(),
                    "descriptors_sha256": hashlib.sha256(d.tobytes()).hexdigest(),
                    "pyramid_sha256": hashlib.sha256(b"".join(ex.pyramid_level(l, True).tobytes() for l in range(8))).hexdigest()}
l, r = synth_stereo_pair(1241, 376, 0)
a, b = orbref.RefExtractor(2000), orbref.RefExtractor(2000)
kl, dl = a.extract(l); kr, dr = b.extract(r)
rc, ur, dp, sad, nm = orbref.stereo_match(a, b, kl, dl, kr, dr, 386.1448, np.float32(386.1448) / np.float32(718.856))
hashes["C2_stereo_s0"] = {"rc": rc, "n_matched": nm, "uright_sha256": hashlib.sha256(ur.tobytes()).hexdigest(),
                          "depth_sha256": hashlib.sha256(dp.tobytes()).hexdigest()}
json.dump({"meta": meta, "hashes": hashes}, open("tests/golden/oracle_hashes.json", "w"), indent=1)
print("wrote tests/golden/cv2_primitives.npz", os.path.getsize("tests/golden/cv2_primitives.npz"), "bytes;", meta)
