"""Writes the C3 inputs of ar_orbslam2_b200/shim/c3_bench into a directory (GPU box; for ncu / sanitizer runs of that binary)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ar_orbslam2_b200.extractor import ORBextractor
from ar_orbslam2_b200.synth import SplitMix64, shifted_frame, synth_frame

td = sys.argv[1]
os.makedirs(td, exist_ok=True)
W3, H3 = 752, 480
FX3 = FY3 = np.float32(435.2); CX3, CY3 = np.float32(367.2), np.float32(252.2)
a = synth_frame(W3, H3, 31)
b = shifted_frame(a, 7, 4, 32)
ex = ORBextractor(2000, 1.2, 8, 20, 7, max_width=W3, max_height=H3)
ka, da = ex(a)
kb, db = ex(b)
scale = ex.GetScaleFactors()
n = len(ka)
z = np.array([2 + (SplitMix64(i).next() % 1000) / 100 for i in range(n)], np.float32)
X, Y = (ka["x"] - CX3) * z / FX3 + np.float32(0.05), (ka["y"] - CY3) * z / FY3 + np.float32(0.02)
invz = (np.float32(1) / z).astype(np.float32)
u, v = (FX3 * X * invz + CX3).astype(np.float32), (FY3 * Y * invz + CY3).astype(np.float32)
valid = ((u >= 0) & (u <= W3) & (v >= 0) & (v <= H3) & (np.arange(n) % 17 != 0)).astype(np.uint8)
rng = SplitMix64(9)
in_view = np.array([rng.next() % 10 != 0 for _ in range(n)], np.uint8)
px, py = (ka["x"] + 7).astype(np.float32), (ka["y"] + 4).astype(np.float32)
w = lambda name, arr: np.ascontiguousarray(arr).tofile(os.path.join(td, name))
w("a.raw", a); w("b.raw", b); w("ka.bin", ka); w("da.bin", da); w("kb.bin", kb); w("db.bin", db); w("scale.bin", scale.astype(np.float32))
w("last_valid.bin", valid); w("last_u.bin", u); w("last_v.bin", v); w("last_invz.bin", invz)
w("last_octave.bin", ka["octave"].astype(np.int32)); w("last_angle.bin", ka["angle"].astype(np.float32)); w("last_obs.bin", np.ones(n, np.uint8))
w("last_claimed.bin", np.zeros(n, np.uint8)); w("last_uright.bin", np.full(n, -1, np.float32))
w("pts_in_view.bin", in_view); w("pts_x.bin", px); w("pts_y.bin", py); w("pts_xr.bin", (px - 6).astype(np.float32)); w("pts_level.bin", ka["octave"].astype(np.int32))
w("pts_view_cos.bin", np.full(n, 0.9, np.float32)); w("pts_obs.bin", np.ones(n, np.uint8)); w("pts_claimed.bin", np.zeros(len(kb), np.uint8))
print("wrote", td)
