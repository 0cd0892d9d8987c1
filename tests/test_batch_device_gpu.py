"""GPU parity of the device-resident batched API that bench.py times (orb_extract_batch with device inputs/outputs,
orb_stereo_match_batch with device arrays): every frame / pair of the batch equals the oracle."""
import numpy as np
import pytest

from ar_orbslam2_b200._abi import KEYPOINT_DTYPE
from ar_orbslam2_b200.synth import synth_stereo_pair

pytestmark = pytest.mark.gpu
MBF, FX = 386.1448, 718.856


def test_stereo_batch_device_resident(oracle, gpu_lib):
    import torch
    from ar_orbslam2_b200.extractor import ORBextractor
    from ar_orbslam2_b200.stereo import stereo_match_batch_device
    w, h, nf, B = 1241, 376, 2000, 3
    pairs = [synth_stereo_pair(w, h, 50 + i) for i in range(B)]
    frames = np.stack([p[0] for p in pairs] + [p[1] for p in pairs])            # lefts then rights, as bench.py lays them out
    dev = torch.device("cuda", 0)
    d_img = torch.from_numpy(frames).to(dev)
    ex = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=2 * B, device=0)
    cap = ex.default_capacity
    d_k = torch.zeros((2 * B, cap, 28), dtype=torch.uint8, device=dev)
    d_d = torch.zeros((2 * B, cap, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(2 * B, dtype=torch.int32, device=dev)
    d_ur = torch.zeros((B, cap), dtype=torch.float32, device=dev)
    d_dp = torch.zeros((B, cap), dtype=torch.float32, device=dev)
    d_nm = torch.zeros(B, dtype=torch.int32, device=dev)
    mb = float(np.float32(MBF) / np.float32(FX))
    for _ in range(2):   # twice: buffers are reused across steps
        ex.extract_batch_device(d_img.data_ptr(), 2 * B, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), cap, d_n.data_ptr())
        stereo_match_batch_device(ex, 0, ex, B, B, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(), d_k[B:].data_ptr(), d_d[B:].data_ptr(),
                                  d_n[B:].data_ptr(), cap, MBF, mb, d_ur.data_ptr(), d_dp.data_ptr(), d_nm.data_ptr())
        ex.sync()
        ex.check()
    n = d_n.cpu().numpy()
    K = d_k.cpu().numpy().reshape(2 * B, cap * 28)
    D = d_d.cpu().numpy()
    UR, DP, NM = d_ur.cpu().numpy(), d_dp.cpu().numpy(), d_nm.cpu().numpy()
    for b in range(B):
        rl, rr = oracle.RefExtractor(nf), oracle.RefExtractor(nf)
        kl, dl = rl.extract(pairs[b][0])
        kr, dr = rr.extract(pairs[b][1])
        assert n[b] == len(kl) and n[B + b] == len(kr)
        assert K[b, : 28 * len(kl)].tobytes() == kl.tobytes() and K[B + b, : 28 * len(kr)].tobytes() == kr.tobytes()
        assert np.array_equal(D[b, : len(kl)], dl) and np.array_equal(D[B + b, : len(kr)], dr)
        rc, ur, dp, _, nm = oracle.stereo_match(rl, rr, kl, dl, kr, dr, MBF, np.float32(MBF) / np.float32(FX))
        assert rc == 0 and nm == NM[b]
        assert np.array_equal(UR[b, : len(kl)].view(np.uint32), ur.view(np.uint32))
        assert np.array_equal(DP[b, : len(kl)].view(np.uint32), dp.view(np.uint32))
        assert (UR[b, len(kl):] == -1).all()
    ex.close()


def test_device_output_overflow_is_latched(gpu_lib):
    """Asynchronous (device-output) batches cannot return an error code; capacity overflow must surface in check()."""
    import torch
    from ar_orbslam2_b200 import _lib
    from ar_orbslam2_b200.extractor import ORBextractor
    from ar_orbslam2_b200.synth import synth_frame
    w, h = 640, 480
    dev = torch.device("cuda", 0)
    d_img = torch.from_numpy(synth_frame(w, h, 1)[None]).to(dev)
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=1, device=0)
    cap = 100
    d_k = torch.zeros((1, cap, 28), dtype=torch.uint8, device=dev)
    d_d = torch.zeros((1, cap, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(1, dtype=torch.int32, device=dev)
    ex.extract_batch_device(d_img.data_ptr(), 1, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), cap, d_n.data_ptr())
    with pytest.raises(_lib.OrbGpuError) as e:
        ex.check()
    assert e.value.code == -4 and int(d_n.item()) > cap      # n_out still reports the required size
    ex.check()                                                # latched flag is cleared by the first check
    ex.close()
