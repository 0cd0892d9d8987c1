"""Shared inputs for the ComputeDistinctiveDescriptors tests: groups of observations of one physical point = one base
descriptor with a few bits flipped per observation, plus outliers, duplicates and degenerate group sizes."""
import numpy as np


def mappoint_groups(seed=0, npoints=300, max_obs=40):
    rng = np.random.default_rng(seed)
    sizes = rng.integers(1, max_obs + 1, npoints)
    sizes[:8] = [1, 2, 3, 4, 31, 32, 33, 64]          # parity of N matters for the median index; warp-size edges
    sizes[8] = 0                                       # no observation left (all key frames bad)
    sizes[9] = 200                                     # a point seen from many key frames: several rows per lane
    rows = []
    for n in sizes:
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        g = np.repeat(base[None], n, axis=0)
        bits = np.unpackbits(g, axis=1)
        for i in range(n):
            k = int(rng.integers(0, 60))
            bits[i, rng.choice(256, k, replace=False)] ^= 1
        g = np.packbits(bits, axis=1)
        if n > 4:
            g[rng.integers(0, n)] = rng.integers(0, 256, 32, dtype=np.uint8)   # an outlier observation
            g[1] = g[0]                                                          # duplicates -> equal medians, first wins
        rows.append(g)
    off = np.zeros(npoints + 1, np.int32)
    off[1:] = np.cumsum(sizes)
    return np.ascontiguousarray(np.concatenate(rows)), off
