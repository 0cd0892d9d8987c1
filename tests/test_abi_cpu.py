"""CPU tests (no GPU): the C-ABI library builds, loads, exports every symbol include/orbgpu.h declares, and refuses
to compute without a device (there is no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from ar_orbslam2_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    return _lib.lib()


def declared_functions():
    src = open(os.path.join(ROOT, "include", "orbgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(orb_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(lib):
    names = declared_functions()
    assert len(names) >= 35
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/orbgpu.h but not exported: {missing}"


def test_abi_version_and_struct_sizes(lib):
    from ar_orbslam2_b200._abi import CANDIDATE_DTYPE, KEYPOINT_DTYPE, OrbParams
    assert lib.orb_abi_version() == 1
    assert KEYPOINT_DTYPE.itemsize == 28 and CANDIDATE_DTYPE.itemsize == 8 and C.sizeof(OrbParams) == 40


def test_built_for_sm_100a():
    out = subprocess.run(["cuobjdump", "-lelf", os.path.join(ROOT, "ar_orbslam2_b200", "liborbgpu.so")], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in out.stdout


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from ar_orbslam2_b200 import _lib
    from ar_orbslam2_b200.extractor import ORBextractor
    from ar_orbslam2_b200.matcher import DescriptorDB, hamming_top2
    with pytest.raises(_lib.OrbGpuError) as e:
        ORBextractor(1000, 1.2, 8, 20, 7)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)
    with pytest.raises(_lib.OrbGpuError):
        DescriptorDB(0, 10)
    with pytest.raises(_lib.OrbGpuError):
        hamming_top2(np.zeros((1, 32), np.uint8), np.zeros((4, 32), np.uint8))
    from ar_orbslam2_b200.synth import synth_vocabulary
    from ar_orbslam2_b200.vocabulary import ORBVocabulary
    from ar_orbslam2_b200 import cvorb
    from ar_orbslam2_b200.stereo import ComputeStereoFromRGBD
    from ar_orbslam2_b200._abi import KEYPOINT_DTYPE
    with pytest.raises(_lib.OrbGpuError) as e:
        cvorb.ORB()
    assert e.value.code == -2
    with pytest.raises(_lib.OrbGpuError):
        cvorb.retain_best(np.arange(10, dtype=np.float32), 3)
    with pytest.raises(_lib.OrbGpuError):
        ComputeStereoFromRGBD(np.zeros(4, KEYPOINT_DTYPE), np.zeros(4, KEYPOINT_DTYPE), np.ones((8, 8), np.float32), 40.0)
    voc = synth_vocabulary(3, 2, seed=0)
    with pytest.raises(_lib.OrbGpuError) as e:
        ORBVocabulary.from_arrays(3, 2, voc["parent"], voc["descriptors"], voc["weight"], voc["word_id"])
    assert e.value.code == -2


def test_product_path_does_not_touch_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may use oracle/."""
    pkg = os.path.join(ROOT, "ar_orbslam2_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "orbref" not in text and "from oracle" not in text and "import oracle" not in text, os.path.join(dirpath, f)
    assert "orbref" not in open(os.path.join(ROOT, "include", "orbgpu.h")).read()
