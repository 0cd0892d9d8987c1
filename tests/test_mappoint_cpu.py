"""MapPoint::ComputeDistinctiveDescriptors: the restated oracle against the reference's own MapPoint.cc
compiled unmodified (oracle/_ref/libmappoint_ref.so)."""
import numpy as np
import pytest

from mappoint_cases import mappoint_groups
from oracle import orbref

needs_ref = pytest.mark.skipif(not orbref.ref_mappoint_available(), reason="oracle/_ref/libmappoint_ref.so not built and no reference checkout")


@needs_ref
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_oracle_equals_reference_mappoint(seed):
    desc, off = mappoint_groups(seed)
    row, med = orbref.distinctive_descriptors(desc, off)
    want = orbref.ref_distinctive_descriptors(desc, off)
    for p in range(len(off) - 1):
        if off[p + 1] == off[p]:
            assert row[p] == -1 and not want[p].any()      # the reference returns early and leaves mDescriptor empty
        else:
            assert np.array_equal(desc[off[p] + row[p]], want[p]), p


def test_median_definition():
    desc, off = mappoint_groups(3, npoints=40)
    row, med = orbref.distinctive_descriptors(desc, off)
    for p in range(len(off) - 1):
        g = desc[off[p]:off[p + 1]]
        n = len(g)
        if n == 0:
            continue
        D = np.unpackbits(g[:, None, :] ^ g[None, :, :], axis=2).sum(axis=2)
        m = np.sort(D, axis=1)[:, (n - 1) // 2]
        assert med[p] == m.min() and row[p] == int(np.argmin(m))
