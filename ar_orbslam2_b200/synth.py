"""Deterministic synthetic inputs: textured frames, stereo pairs,
descriptor databases.  PRNG = splitmix64, so the same frames can be regenerated anywhere
(numpy here; the C++ side only ever receives the pixels).

The reference ships no dataset (TUM/KITTI/EuRoC are external downloads, ORB_SLAM2/README.md),
so every benchmark and parity test runs on these.
"""
from __future__ import annotations

import numpy as np

_GAMMA = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)
_MASK = (1 << 64) - 1


def _mix(z: np.ndarray) -> np.ndarray:
    """splitmix64 output function on a uint64 array (wrapping arithmetic)."""
    with np.errstate(over="ignore"):
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


class SplitMix64:
    """Scalar splitmix64 stream (python ints, exact)."""

    def __init__(self, seed: int):
        self.state = seed & _MASK

    def next(self) -> int:
        self.state = (self.state + 0x9E3779B97F4A7C15) & _MASK
        z = self.state
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _MASK
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _MASK
        return z ^ (z >> 31)


def counter_stream(seed: int, n: int) -> np.ndarray:
    """n splitmix64 outputs of the stream seeded with `seed`, vectorised (counter form)."""
    with np.errstate(over="ignore"):
        idx = np.arange(1, n + 1, dtype=np.uint64)
        return _mix(np.uint64(seed & _MASK) + idx * _GAMMA)


def synth_frame(width: int, height: int, seed: int) -> np.ndarray:
    """Texture frame: grey-128 canvas, w*h/600 filled rectangles (4..59 px sides, random grey),
    uniform integer noise in [-6, 6], clipped to u8.  ~8k FAST candidates at 640x480."""
    rng = SplitMix64(0xA5A5_0000_0000_0000 ^ (seed * 0x1000193 + 1))
    img = np.full((height, width), 128, dtype=np.int16)
    for _ in range(width * height // 600):
        x0 = rng.next() % width
        y0 = rng.next() % height
        rw = 4 + rng.next() % 56
        rh = 4 + rng.next() % 56
        g = rng.next() % 256
        img[y0:min(y0 + rh, height), x0:min(x0 + rw, width)] = g
    noise = counter_stream(0x5EED_0000_0000_0000 ^ (seed * 0x9E37 + 7), width * height)
    img += (noise % np.uint64(13)).astype(np.int16).reshape(height, width) - 6
    return np.clip(img, 0, 255).astype(np.uint8)


def synth_stereo_pair(width: int, height: int, seed: int, band_rows: int = 47):
    """Left = synth_frame; right = left shifted left by a band-wise constant disparity
    d in {4, 8, ..., 64} px per `band_rows`-row band; the uncovered strip on the right edge is
    filled from a second seed."""
    left = synth_frame(width, height, seed)
    filler = synth_frame(width, height, seed + 1_000_003)
    right = np.empty_like(left)
    nb = (height + band_rows - 1) // band_rows
    for b in range(nb):
        d = 4 * (1 + (b * 7 + seed) % 16)
        r0, r1 = b * band_rows, min((b + 1) * band_rows, height)
        right[r0:r1, : width - d] = left[r0:r1, d:]
        right[r0:r1, width - d:] = filler[r0:r1, width - d:]
    return left, right


def shifted_frame(frame: np.ndarray, dx: int, dy: int, seed: int) -> np.ndarray:
    """frame translated by (+dx,+dy) pixels (content moves right/down); uncovered area from `seed`."""
    h, w = frame.shape
    out = synth_frame(w, h, seed)
    out[max(dy, 0):h + min(dy, 0), max(dx, 0):w + min(dx, 0)] = frame[max(-dy, 0):h - max(dy, 0), max(-dx, 0):w - max(dx, 0)]
    return out


def synth_descriptors(n: int, seed: int, first_row: int = 0) -> np.ndarray:
    """n x 32 uniform random descriptor bytes; row r depends only on (seed, global row index),
    so any shard of the database can be generated independently (config C5).  Matches the
    device-side generator orb_db_fill_synthetic."""
    with np.errstate(over="ignore"):
        words = np.arange(first_row * 4, (first_row + n) * 4, dtype=np.uint64)
        z = _mix(np.uint64(seed & _MASK) + (words + np.uint64(1)) * _GAMMA)
    return z.view(np.uint8).reshape(n, 32).copy()


def synth_queries(db_rows_fn, n_db: int, nq: int, seed: int) -> np.ndarray:
    """nq query descriptors: even rows are database rows with 0..40 random bit flips (so the
    TH_LOW / ratio paths are exercised), odd rows are random.  `db_rows_fn(idx_array)` returns the
    database rows for the given global indices."""
    rng = SplitMix64(0xC0FFEE ^ seed)
    q = synth_descriptors(nq, seed ^ 0xABCDEF)
    pick = np.array([rng.next() % n_db for _ in range((nq + 1) // 2)], dtype=np.int64)
    rows = db_rows_fn(pick).copy()
    for i in range(rows.shape[0]):
        for _ in range(rng.next() % 41):
            bit = rng.next() % 256
            rows[i, bit >> 3] ^= np.uint8(1 << (bit & 7))
    q[0::2] = rows
    return q
