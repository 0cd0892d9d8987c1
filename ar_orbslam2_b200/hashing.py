"""SHA-256 of extraction / stereo results in the layout tools/make_golden_sweeps.py hashes the oracle's outputs in, so a
GPU box can check whole batches against tests/golden/oracle_sweep_hashes.json without running the oracle."""
import hashlib

import numpy as np


def frame_hash(keypoints: np.ndarray, descriptors: np.ndarray, n: int) -> str:
    """keypoints: [>= n] 28-byte records (any dtype of 28 bytes per row), descriptors: [>= n][32] u8."""
    k = np.ascontiguousarray(keypoints).view(np.uint8).reshape(-1)[: 28 * n]
    d = np.ascontiguousarray(descriptors).view(np.uint8).reshape(-1)[: 32 * n]
    return hashlib.sha256(k.tobytes() + d.tobytes()).hexdigest()


def stereo_hash(u_right: np.ndarray, depth: np.ndarray, n_left: int) -> str:
    return hashlib.sha256(np.ascontiguousarray(u_right[:n_left], np.float32).tobytes() +
                          np.ascontiguousarray(depth[:n_left], np.float32).tobytes()).hexdigest()


def check_stereo_batch(golden_pairs, n, kps, desc, u_right, depth, n_matched):
    """Device-resident batch laid out as bench.py does (B lefts then B rights; row b of the stereo arrays = pair b), already
    copied to host arrays.  golden_pairs[i] is the golden record of pair b = i (mod len).  Returns the list of mismatches."""
    B = len(n) // 2
    bad = []
    for b in range(B):
        g = golden_pairs[b % len(golden_pairs)]
        for side, f in (("left", b), ("right", B + b)):
            if int(n[f]) != g[side]["n"] or frame_hash(kps[f], desc[f], int(n[f])) != g[side]["sha256"]:
                bad.append((b, side))
        if int(n_matched[b]) != g["stereo"]["n_matched"] or stereo_hash(u_right[b], depth[b], int(n[b])) != g["stereo"]["sha256"]:
            bad.append((b, "stereo"))
    return bad
