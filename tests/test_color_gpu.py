"""GPU parity of orb_cvt_color_gray through the C ABI: bit-exact against cv2's golden outputs and the oracle."""
import numpy as np
import pytest

from ar_orbslam2_b200.color import cvtColorGray, cvtColorGray_device
from oracle import orbref
from test_color_cpu import golden_cases, image

pytestmark = pytest.mark.gpu


def test_gpu_equals_cv2_golden():
    for key, img, rgb, want in golden_cases():
        assert np.array_equal(cvtColorGray(img, rgb), want), key


@pytest.mark.parametrize("h,w,ch", [(376, 1241, 3), (480, 640, 4), (1, 1, 3), (3, 15, 4), (7, 4097, 3), (1080, 1920, 3)])
def test_gpu_equals_oracle(h, w, ch):
    img = image(h * 7 + w, h, w, ch)
    for rgb in (True, False):
        for variant in (0, 1):
            assert np.array_equal(cvtColorGray(img, rgb, variant), orbref.cvt_color_gray(img, rgb, variant))


def test_batched_device_resident_and_feeds_the_extractor():
    import torch
    from ar_orbslam2_b200.extractor import ORBextractor
    from ar_orbslam2_b200.synth import synth_frame
    dev = torch.device("cuda:0")
    w, h, B = 640, 480, 3
    grey = np.stack([synth_frame(w, h, 10 + i) for i in range(B)])
    rng = np.random.default_rng(0)
    # colour frames whose grey conversion is known: R=G=B=grey converts to grey exactly (9798+19235+3735 = 32768)
    rgb = np.repeat(grey[..., None], 3, axis=3)
    rgb[..., 0] = np.clip(rgb[..., 0].astype(int) + rng.integers(-3, 4, grey.shape), 0, 255)
    want = np.stack([orbref.cvt_color_gray(rgb[b], True) for b in range(B)])
    d_rgb = torch.from_numpy(np.ascontiguousarray(rgb)).to(dev)
    d_grey = torch.zeros((B, h, w), dtype=torch.uint8, device=dev)
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B, device=0)
    cvtColorGray_device(d_rgb.data_ptr(), w, h, w * 3, h * w * 3, 3, True, B, d_grey.data_ptr(), w, h * w, stream=ex.stream)
    cap = ex.default_capacity
    d_k = torch.zeros((B, cap, 28), dtype=torch.uint8, device=dev)
    d_d = torch.zeros((B, cap, 32), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(B, dtype=torch.int32, device=dev)
    ex.extract_batch_device(d_grey.data_ptr(), B, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), cap, d_n.data_ptr())
    ex.check()
    assert np.array_equal(d_grey.cpu().numpy(), want)
    ref = orbref.RefExtractor(1000)
    for b in range(B):
        rk, rd = ref.extract(want[b])
        n = int(d_n[b])
        assert n == len(rk) and np.array_equal(d_d[b, :n].cpu().numpy(), rd)
    ex.close()
