#!/usr/bin/env python3
"""Generate tests/golden/oracle_sweep_hashes.json: SHA-256 of the CPU oracle's full outputs (oracle/orbref_*.cpp through
oracle/orbref.py) on seed sweeps of the benchmark configurations, so that the GPU box only has to hash its own outputs.

  C1      640x480, 1000 features, seeds 0..99           one hash per frame over key points + descriptors
  C2      1241x376 stereo pairs, 2000 features, 0..31   left / right frame hashes + hash over mvuRight + mvDepth
  C4      1920x1080, 2000 features, seeds 0..15
  bench   the pairs bench.py tiles into its 288-frame step: seeds 17 + 1000 * rank + i, rank 0..7, i 0..15

Runs on the CPU box (the oracle takes ~15 ms per VGA frame); usage: python tools/make_golden_sweeps.py"""
import hashlib
import json
import os
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ar_orbslam2_b200.synth import synth_frame, synth_stereo_pair   # noqa: E402
from oracle import orbref                                           # noqa: E402

MBF, FX = 386.1448, 718.856


def frame_hash(k, d):
    return hashlib.sha256(k.tobytes() + d.tobytes()).hexdigest()


def one_frame(args):
    w, h, nf, seed = args
    k, d = orbref.RefExtractor(nf).extract(synth_frame(w, h, seed))
    return {"n": int(len(k)), "sha256": frame_hash(k, d)}


def one_pair(args):
    w, h, nf, seed = args
    l, r = synth_stereo_pair(w, h, seed)
    a, b = orbref.RefExtractor(nf), orbref.RefExtractor(nf)
    kl, dl = a.extract(l)
    kr, dr = b.extract(r)
    rc, ur, dp, _, nm = orbref.stereo_match(a, b, kl, dl, kr, dr, MBF, np.float32(MBF) / np.float32(FX))
    return {"left": {"n": int(len(kl)), "sha256": frame_hash(kl, dl)}, "right": {"n": int(len(kr)), "sha256": frame_hash(kr, dr)},
            "stereo": {"rc": int(rc), "n_matched": int(nm), "sha256": hashlib.sha256(ur.tobytes() + dp.tobytes()).hexdigest()}}


def main():
    orbref.build()
    with ThreadPoolExecutor(os.cpu_count() or 1) as pool:
        out = {
            "meta": {"generator": "tools/make_golden_sweeps.py", "frame_hash": "sha256(keypoints[n] as 28-byte records + descriptors[n][32])",
                     "stereo_hash": "sha256(mvuRight[n_left] f32 + mvDepth[n_left] f32)", "mbf": MBF, "fx": FX},
            "C1": list(pool.map(one_frame, [(640, 480, 1000, s) for s in range(100)])),
            "C2": list(pool.map(one_pair, [(1241, 376, 2000, s) for s in range(32)])),
            "C4": list(pool.map(one_frame, [(1920, 1080, 2000, s) for s in range(16)])),
            "bench": {str(rank): list(pool.map(one_pair, [(1241, 376, 2000, 17 + 1000 * rank + i) for i in range(16)])) for rank in range(8)},
        }
    path = os.path.join(ROOT, "tests", "golden", "oracle_sweep_hashes.json")
    json.dump(out, open(path, "w"), indent=0)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
