"""GPU parity of orb_distinctive_descriptors (MapPoint::ComputeDistinctiveDescriptors) through the C ABI."""
import os

import numpy as np
import pytest

from ar_orbslam2_b200.matcher import ComputeDistinctiveDescriptors
from mappoint_cases import mappoint_groups
from oracle import orbref

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_gpu_equals_oracle_and_reference(seed):
    desc, off = mappoint_groups(seed)
    row, med = ComputeDistinctiveDescriptors(desc, off)
    orow, omed = orbref.distinctive_descriptors(desc, off)
    assert np.array_equal(row, orow) and np.array_equal(med, omed)
    if os.path.exists(orbref.REF_MAPPOINT_PATH):
        want = orbref.ref_distinctive_descriptors(desc, off)
        for p in range(len(off) - 1):
            if off[p + 1] > off[p]:
                assert np.array_equal(desc[off[p] + row[p]], want[p]), p


def test_large_batch_property():
    """20 000 map points: the chosen row's median is the minimum over its group (checked in numpy on a sample)."""
    desc, off = mappoint_groups(9, npoints=20000, max_obs=12)
    row, med = ComputeDistinctiveDescriptors(desc, off)
    orow, omed = orbref.distinctive_descriptors(desc, off)
    assert np.array_equal(row, orow) and np.array_equal(med, omed)


def test_empty_call():
    row, med = ComputeDistinctiveDescriptors(np.zeros((0, 32), np.uint8), np.zeros(1, np.int32))
    assert len(row) == 0
