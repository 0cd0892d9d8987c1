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


# ---- synthetic bag-of-words vocabulary (the reference's ORBvoc.txt is an external download, ORB_SLAM2/README.md) -----

def synth_vocabulary(k: int = 10, L: int = 3, seed: int = 0, flip_bits: int = 40, stop_fraction: float = 0.02,
                     dup_fraction: float = 0.05, early_leaf_fraction: float = 0.0):
    """A full k-ary tree of depth L in the node order DBoW2's create()/loadFromTextFile produce (a node's id is larger
    than its parent's; children of a node in ascending id).  A child's descriptor is its parent's with `flip_bits`
    random bit flips, so the descent is decided by real distances; `dup_fraction` of the children copy their left
    sibling's descriptor (distance ties -> first child must win); `stop_fraction` of the words have weight 0
    ("stopped" words); `early_leaf_fraction` of the inner nodes below the root are cut to leaves.
    Returns dict(k, L, parent, descriptors, weight, word_id, is_leaf) with node 0 = root."""
    rng = np.random.default_rng(seed)
    parent, desc, depth = [0], [np.zeros(32, np.uint8)], [0]
    frontier = [0]
    root_desc = rng.integers(0, 256, 32, dtype=np.uint8)
    for lvl in range(1, L + 1):
        nxt = []
        for p in frontier:
            if lvl > 1 and lvl <= L and early_leaf_fraction > 0 and rng.random() < early_leaf_fraction:
                continue   # p stays a leaf
            base = root_desc if p == 0 else desc[p]
            prev = None
            for c in range(k):
                if prev is not None and rng.random() < dup_fraction:
                    d = prev.copy()
                else:
                    bits = np.unpackbits(base.copy())
                    flips = rng.choice(256, size=flip_bits, replace=False)
                    bits[flips] ^= 1
                    d = np.packbits(bits)
                prev = d
                parent.append(p); desc.append(d); depth.append(lvl)
                nxt.append(len(parent) - 1)
        frontier = nxt
    n = len(parent)
    parent = np.asarray(parent, np.int32)
    desc = np.stack(desc).astype(np.uint8)
    has_child = np.zeros(n, bool)
    has_child[parent[1:]] = True
    is_leaf = ~has_child
    is_leaf[0] = False
    word_id = np.zeros(n, np.int32)
    word_id[is_leaf] = np.arange(int(is_leaf.sum()), dtype=np.int32)      # word ids in node order (loadFromTextFile :1409-1416)
    weight = np.zeros(n, np.float64)
    w = rng.uniform(0.5, 12.0, int(is_leaf.sum()))                         # idf-like values
    w[rng.random(w.shape[0]) < stop_fraction] = 0.0
    weight[is_leaf] = w
    return dict(k=k, L=L, parent=parent, descriptors=desc, weight=weight, word_id=word_id, is_leaf=is_leaf)


def write_vocabulary_text(path, voc, scoring: int = 0, weighting: int = 0, trailing_newline: bool = True):
    """ORBvoc.txt format (TemplatedVocabulary::saveToTextFile, TemplatedVocabulary.h:1428-1448): header `k L  scoring
    weighting`, then per node `parent is_leaf d0 .. d31 weight`."""
    lines = [f"{voc['k']} {voc['L']}  {scoring} {weighting}"]
    for i in range(1, voc["parent"].shape[0]):
        d = " ".join(str(int(x)) for x in voc["descriptors"][i])
        lines.append(f"{int(voc['parent'][i])} {1 if voc['is_leaf'][i] else 0} {d}  {repr(float(voc['weight'][i]))}")
    with open(path, "w") as f:
        f.write("\n".join(lines) + ("\n" if trailing_newline else ""))


def synth_bow_descriptors(voc, n: int, seed: int = 0, flip_bits: int = 30) -> np.ndarray:
    """n query descriptors: random words of the vocabulary with `flip_bits` bits flipped (so they descend somewhere
    meaningful), a few exact copies of inner nodes and a few uniformly random rows."""
    rng = np.random.default_rng(seed + 7919)
    leaves = np.nonzero(voc["is_leaf"])[0]
    pick = rng.choice(leaves, size=n)
    out = voc["descriptors"][pick].copy()
    bits = np.unpackbits(out, axis=1)
    for i in range(n):
        bits[i, rng.choice(256, size=flip_bits, replace=False)] ^= 1
    out = np.packbits(bits, axis=1)
    nr = max(1, n // 16)
    out[rng.choice(n, nr, replace=False)] = rng.integers(0, 256, (nr, 32), dtype=np.uint8)
    return np.ascontiguousarray(out)


def synth_vocabulary_fast(k: int = 10, L: int = 6, seed: int = 0, flip_prob: float = 0.15):
    """ORBvoc-sized trees (k=10, L=6: 1.1 M nodes) in seconds: level by level, a child's descriptor = its parent's XOR a
    random mask with each bit set with probability `flip_prob`.  Same node order and fields as synth_vocabulary."""
    rng = np.random.default_rng(seed)
    parents, descs = [np.zeros(1, np.int32)], [rng.integers(0, 256, (1, 32), dtype=np.uint8)]
    first = 0
    for lvl in range(1, L + 1):
        prev = descs[-1]
        n_prev = prev.shape[0]
        par = np.repeat(np.arange(first, first + n_prev, dtype=np.int32), k)
        child = np.repeat(prev, k, axis=0)
        for s in range(0, child.shape[0], 1 << 17):
            e = min(child.shape[0], s + (1 << 17))
            mask = np.packbits(rng.random((e - s, 256), dtype=np.float32) < flip_prob, axis=1)
            child[s:e] ^= mask
        first += n_prev
        parents.append(par)
        descs.append(child)
    parent = np.concatenate(parents)
    desc = np.concatenate(descs)
    desc[0] = 0
    n = parent.shape[0]
    n_leaf = descs[-1].shape[0]
    is_leaf = np.zeros(n, bool)
    is_leaf[n - n_leaf:] = True
    word_id = np.zeros(n, np.int32)
    word_id[is_leaf] = np.arange(n_leaf, dtype=np.int32)
    weight = np.zeros(n, np.float64)
    weight[is_leaf] = rng.uniform(0.5, 12.0, n_leaf)
    return dict(k=k, L=L, parent=parent, descriptors=desc, weight=weight, word_id=word_id, is_leaf=is_leaf)
