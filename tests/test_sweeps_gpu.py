"""Seed sweeps of the benchmark configurations as ONE batched GPU call each, checked against SHA-256 hashes of the CPU
oracle's outputs (tests/golden/oracle_sweep_hashes.json, written by tools/make_golden_sweeps.py on the CPU box; the oracle
itself is pinned to the reference's own sources by tests/test_reference_cpu.py).  The GPU box only hashes.
ref: ORB_SLAM2/src/ORBextractor.cc:983-1047, ORB_SLAM2/src/Frame.cc:471-643."""
import json
import os

import numpy as np
import pytest

from ar_orbslam2_b200.hashing import check_stereo_batch, frame_hash
from ar_orbslam2_b200.synth import synth_frame, synth_stereo_pair

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
MBF, FX = 386.1448, 718.856


@pytest.fixture(scope="module")
def golden():
    return json.load(open(os.path.join(HERE, "golden", "oracle_sweep_hashes.json")))


def _device_batch(frames, nf, w, h):
    import torch
    from ar_orbslam2_b200.extractor import ORBextractor
    nb = len(frames)
    dev = torch.device("cuda", 0)
    ex = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=nb, device=0)
    cap = ex.default_capacity
    d_img = torch.from_numpy(np.stack(frames)).to(dev)
    d_k = torch.zeros((nb, cap, 28), dtype=torch.uint8, device=dev)
    d_d = torch.zeros((nb, cap, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(nb, dtype=torch.int32, device=dev)
    ex.extract_batch_device(d_img.data_ptr(), nb, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), cap, d_n.data_ptr())
    ex.sync()
    ex.check()
    return ex, cap, d_k, d_d, d_n


@pytest.mark.parametrize("cfg,w,h,nf,nseeds", [("C1", 640, 480, 1000, 100), ("C4", 1920, 1080, 2000, 16)])
def test_mono_seed_sweep_one_batched_call(gpu_lib, golden, cfg, w, h, nf, nseeds):
    frames = [synth_frame(w, h, s) for s in range(nseeds)]
    ex, cap, d_k, d_d, d_n = _device_batch(frames, nf, w, h)
    n, K, D = d_n.cpu().numpy(), d_k.cpu().numpy(), d_d.cpu().numpy()
    bad = [s for s in range(nseeds) if int(n[s]) != golden[cfg][s]["n"] or frame_hash(K[s], D[s], int(n[s])) != golden[cfg][s]["sha256"]]
    assert not bad, f"{cfg}: seeds {bad} differ from the oracle"
    ex.close()


def test_c2_stereo_seed_sweep_one_batched_call(gpu_lib, golden):
    import torch
    from ar_orbslam2_b200.stereo import stereo_match_batch_device
    w, h, nf, B = 1241, 376, 2000, 32
    pairs = [synth_stereo_pair(w, h, s) for s in range(B)]
    ex, cap, d_k, d_d, d_n = _device_batch([p[0] for p in pairs] + [p[1] for p in pairs], nf, w, h)
    dev = d_k.device
    d_ur = torch.zeros((B, cap), dtype=torch.float32, device=dev)
    d_dp = torch.zeros((B, cap), dtype=torch.float32, device=dev)
    d_nm = torch.zeros(B, dtype=torch.int32, device=dev)
    mb = float(np.float32(MBF) / np.float32(FX))
    stereo_match_batch_device(ex, 0, ex, B, B, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), d_k[B:].data_ptr(), d_d[B:].data_ptr(),
                              d_n[B:].data_ptr(), cap, MBF, mb, d_ur.data_ptr(), d_dp.data_ptr(), d_nm.data_ptr())
    ex.sync()
    ex.check()
    bad = check_stereo_batch(golden["C2"], d_n.cpu().numpy(), d_k.cpu().numpy(), d_d.cpu().numpy(), d_ur.cpu().numpy(),
                             d_dp.cpu().numpy(), d_nm.cpu().numpy())
    assert not bad, f"C2: (pair, part) {bad} differ from the oracle"
    ex.close()


def test_bench_step_outputs_by_hash(gpu_lib, golden):
    """The 288-frame step bench.py times (144 pairs tiled from 16 distinct ones, rank 0's seeds): every frame and every pair."""
    import torch
    from ar_orbslam2_b200.stereo import stereo_match_batch_device
    w, h, nf, B = 1241, 376, 2000, 144
    distinct = [synth_stereo_pair(w, h, 17 + i) for i in range(16)]
    frames = [distinct[b % 16][0] for b in range(B)] + [distinct[b % 16][1] for b in range(B)]
    ex, cap, d_k, d_d, d_n = _device_batch(frames, nf, w, h)
    dev = d_k.device
    d_ur = torch.zeros((B, cap), dtype=torch.float32, device=dev)
    d_dp = torch.zeros((B, cap), dtype=torch.float32, device=dev)
    d_nm = torch.zeros(B, dtype=torch.int32, device=dev)
    mb = float(np.float32(MBF) / np.float32(FX))
    stereo_match_batch_device(ex, 0, ex, B, B, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), d_k[B:].data_ptr(), d_d[B:].data_ptr(),
                              d_n[B:].data_ptr(), cap, MBF, mb, d_ur.data_ptr(), d_dp.data_ptr(), d_nm.data_ptr())
    ex.sync()
    ex.check()
    bad = check_stereo_batch(golden["bench"]["0"], d_n.cpu().numpy(), d_k.cpu().numpy(), d_d.cpu().numpy(), d_ur.cpu().numpy(),
                             d_dp.cpu().numpy(), d_nm.cpu().numpy())
    assert not bad, f"bench step: (pair, part) {bad} differ from the oracle"
    ex.close()
