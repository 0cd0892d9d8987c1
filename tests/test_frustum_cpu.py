"""Frame::isInFrustum: the restated oracle against the reference's own Frame.cc compiled unmodified
(oracle/_ref/liborbmatcher_ref.so), every output field bit for bit."""
import numpy as np
import pytest

from frustum_cases import CX, CY, FX, FY, LOGS, MBF, cam_dict, dummy_frame, points, pose
from oracle import orbref

needs_ref = pytest.mark.skipif(not orbref.ref_matcher_available(), reason="oracle/_ref not built and no reference checkout")


def same(a, b):
    for x, y in zip(a, b):
        assert x.dtype == y.dtype and x.tobytes() == y.tobytes()


@needs_ref
@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_oracle_equals_reference_frame_cc(seed):
    f = dummy_frame(orbref.ReferenceSourceBackend())
    wp, nr, maxd, mind, consider = points(seed)
    cam15, want = orbref.refm_is_in_frustum(f, pose(seed), FX, FY, CX, CY, MBF, LOGS, wp, nr, maxd, mind, 0.5, consider)
    got = orbref.is_in_frustum(cam_dict(cam15), wp, nr, maxd, mind, 0.5, consider)
    same(got, want)
    iv = want[0].astype(bool)
    assert 50 < iv.sum() < len(iv) - 50 and len(np.unique(want[4][iv])) > 4      # a spread of predicted levels


def glibc_logf_np(x):
    """numpy restatement of glibc's logf (sysdeps/ieee754/flt-32/e_logf.c) = the algorithm of csrc/orb_frustum.cu::glibc_logf."""
    T = np.array([[float.fromhex('0x1.661ec79f8f3bep+0'), float.fromhex('-0x1.57bf7808caadep-2')], [float.fromhex('0x1.571ed4aaf883dp+0'), float.fromhex('-0x1.2bef0a7c06ddbp-2')],
                  [float.fromhex('0x1.49539f0f010b0p+0'), float.fromhex('-0x1.01eae7f513a67p-2')], [float.fromhex('0x1.3c995b0b80385p+0'), float.fromhex('-0x1.b31d8a68224e9p-3')],
                  [float.fromhex('0x1.30d190c8864a5p+0'), float.fromhex('-0x1.6574f0ac07758p-3')], [float.fromhex('0x1.25e227b0b8ea0p+0'), float.fromhex('-0x1.1aa2bc79c8100p-3')],
                  [float.fromhex('0x1.1bb4a4a1a343fp+0'), float.fromhex('-0x1.a4e76ce8c0e5ep-4')], [float.fromhex('0x1.12358f08ae5bap+0'), float.fromhex('-0x1.1973c5a611cccp-4')],
                  [float.fromhex('0x1.0953f419900a7p+0'), float.fromhex('-0x1.252f438e10c1ep-5')], [1.0, 0.0], [float.fromhex('0x1.e608cfd9a47acp-1'), float.fromhex('0x1.aa5aa5df25984p-5')],
                  [float.fromhex('0x1.ca4b31f026aa0p-1'), float.fromhex('0x1.c5e53aa362eb4p-4')], [float.fromhex('0x1.b2036576afce6p-1'), float.fromhex('0x1.526e57720db08p-3')],
                  [float.fromhex('0x1.9c2d163a1aa2dp-1'), float.fromhex('0x1.bc2860d224770p-3')], [float.fromhex('0x1.886e6037841edp-1'), float.fromhex('0x1.1058bc8a07ee1p-2')],
                  [float.fromhex('0x1.767dcf5534862p-1'), float.fromhex('0x1.4043057b6ee09p-2')]])
    ix = x.view(np.uint32).astype(np.int64)
    tmp = (ix - 0x3f330000) & 0xffffffff
    i = (tmp >> 19) & 15
    k = np.where(tmp >= 1 << 31, tmp - (1 << 32), tmp) >> 23
    iz = (ix - (tmp & (0x1ff << 23))) & 0xffffffff
    z = iz.astype(np.uint32).view(np.float32).astype(np.float64)
    r = z * T[i, 0] - 1
    y0 = T[i, 1] + k.astype(np.float64) * float.fromhex('0x1.62e42fefa39efp-1')
    r2 = r * r
    y = float.fromhex('0x1.5575b0be00b6ap-2') * r + float.fromhex('-0x1.ffffef20a4123p-2')
    y = float.fromhex('-0x1.00ea348b88334p-2') * r2 + y
    y = y * r2 + (y0 + r)
    out = y.astype(np.float32)
    out[ix == 0x3f800000] = 0
    return out


def test_logf_restatement_equals_the_host_libm():
    """The device restates glibc's logf (CUDA's own logf differs in the last bit on some inputs, which would move a predicted
    level).  Exhaustive check: tools/check_logf_exhaustive.c (0 mismatches over all positive finite floats); sampled here."""
    import ctypes as C
    libm = C.CDLL("libm.so.6")
    libm.logf.restype = C.c_float
    libm.logf.argtypes = [C.c_float]
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(0.01, 50, 30000), np.exp(rng.uniform(-80, 80, 30000))]).astype(np.float32)
    host = np.array([libm.logf(float(v)) for v in x], np.float32)
    assert np.array_equal(glibc_logf_np(x).view(np.uint32), host.view(np.uint32))
