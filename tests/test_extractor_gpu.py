"""GPU parity: liborbgpu's extractor (through the C ABI) against the CPU oracle, stage by stage and
end to end, bit-exact.  ref: ORB_SLAM2/src/ORBextractor.cc."""
import numpy as np
import pytest

from ar_orbslam2_b200.synth import synth_frame

pytestmark = pytest.mark.gpu

CONFIGS = [  # (width, height, nfeatures, seed)   SURVEY 8: C1, C2, C3, C1-init, C4
    (640, 480, 1000, 0),
    (1241, 376, 2000, 3),
    (752, 480, 1000, 5),
    (640, 480, 2000, 11),
    (1920, 1080, 2000, 8),
]


def _sorted(c):
    return c


@pytest.mark.parametrize("w,h,nf,seed", CONFIGS)
def test_stages_and_end_to_end(oracle, gpu_lib, w, h, nf, seed):
    from ar_orbslam2_b200.extractor import ORBextractor
    img = synth_frame(w, h, seed)
    ref = oracle.RefExtractor(nf)
    rk, rd = ref.extract(img)
    ex = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    gk, gd = ex(img)
    # E3 pyramid incl. the 19-px reflect-101 frame
    for l in range(8):
        assert ex.level_size(l) == ref.level_size(l)
        a, b = ex.pyramid_level(l, True), ref.pyramid_level(l, True)
        assert np.array_equal(a, b), f"pyramid level {l}: {int((a != b).sum())} pixels differ"
    # E4/E5 FAST candidates in reference order
    for l in range(8):
        a, b = ex.candidates(l), ref.candidates(l)
        assert len(a) == len(b), f"level {l}: {len(a)} candidates vs {len(b)}"
        assert np.array_equal(a, b), f"level {l}: candidate lists differ"
    # E8 blur
    for l in (0, 3, 7):
        assert np.array_equal(ex.blurred_level(l), ref.blurred_level(l)), f"blur level {l}"
    # E6/E7/E9/E10 final keypoints + descriptors, exact order
    assert len(gk) == len(rk), f"{len(gk)} keypoints vs {len(rk)}"
    for f in ("octave", "x", "y", "response", "size", "class_id"):
        assert np.array_equal(gk[f], rk[f]), f"keypoint field {f} differs at {np.flatnonzero(gk[f] != rk[f])[:8]}"
    assert np.array_equal(gk["angle"].view(np.uint32), rk["angle"].view(np.uint32)), "angles differ bitwise"
    assert np.array_equal(gd, rd), f"{int((gd != rd).any(axis=1).sum())} descriptors differ"
    ex.close()


def test_batch_matches_single(oracle, gpu_lib):
    from ar_orbslam2_b200.extractor import ORBextractor
    w, h = 640, 480
    frames = np.stack([synth_frame(w, h, s) for s in (21, 22, 23, 24, 25)])
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=5)
    out = ex.extract_batch(frames)
    ref = oracle.RefExtractor(1000)
    for b in range(5):
        rk, rd = ref.extract(frames[b])
        gk, gd = out[b]
        assert len(gk) == len(rk)
        assert gk.tobytes() == rk.tobytes()
        assert np.array_equal(gd, rd)
    ex.close()


def test_geometry_change_and_errors(oracle, gpu_lib):
    from ar_orbslam2_b200 import _lib
    from ar_orbslam2_b200.extractor import ORBextractor
    ex = ORBextractor(500, 1.2, 8, 20, 7, max_width=800, max_height=600)
    ref = oracle.RefExtractor(500)
    for (w, h, s) in [(800, 600, 1), (320, 240, 2), (641, 479, 3)]:
        img = synth_frame(w, h, s)
        gk, gd = ex(img)
        rk, rd = ref.extract(img)
        assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd), (w, h)
    # empty image: silent, zero keypoints (ref :987-988)
    k, d = ex(np.zeros((0, 0), np.uint8))
    assert len(k) == 0
    # an image whose top level is too small for the 30-px grid: explicit error, as the oracle
    small = synth_frame(160, 120, 4)
    rc, _, _ = ref.extract(small, check=False)
    assert rc == -3
    with pytest.raises(_lib.OrbGpuError) as ei:
        ex(small)
    assert ei.value.code == -3
    # too large for the handle
    with pytest.raises(_lib.OrbGpuError):
        ex(synth_frame(1024, 600, 5))
    # capacity too small is reported, never truncated silently
    with pytest.raises(_lib.OrbGpuError) as ei:
        ex(synth_frame(640, 480, 6), capacity=100)
    assert ei.value.code == -4
    ex.close()


def test_flat_and_noise_images(oracle, gpu_lib):
    """Edge cases: no corners at all; dense noise (many FAST hits, threshold-7 retries)."""
    from ar_orbslam2_b200.extractor import ORBextractor
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480)
    ref = oracle.RefExtractor(1000)
    flat = np.full((480, 640), 77, np.uint8)
    k, d = ex(flat)
    assert len(k) == 0 and d.shape == (0, 32)
    rng = np.random.default_rng(5)
    noise = rng.integers(0, 256, (480, 640), dtype=np.uint8)
    gk, gd = ex(noise)
    rk, rd = ref.extract(noise)
    assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd)
    soft = (128 + 20 * rng.standard_normal((480, 640))).clip(0, 255).astype(np.uint8)
    gk, gd = ex(soft)
    rk, rd = ref.extract(soft)
    assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd)
    # steep wrapping ramps: most pixels are brighter than one compass pair and darker than the other (the both-type list)
    ramp = ((np.arange(640)[None, :] * 9 + np.arange(480)[:, None] * 11) % 256).astype(np.uint8)
    gk, gd = ex(ramp)
    rk, rd = ref.extract(ramp)
    assert len(rk) > 0 and gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd)
    for l in range(8):
        assert np.array_equal(ex.candidates(l), ref.candidates(l)), l
    ex.close()


def test_gpu_equals_reference_source(oracle, gpu_lib):
    """liborbgpu against oracle/_ref (the reference's own ORBextractor.cc, see tests/test_reference_cpu.py)."""
    from ar_orbslam2_b200.extractor import ORBextractor
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built")
    for (w, h, nf, seed) in [(640, 480, 1000, 40), (1241, 376, 2000, 41)]:
        img = synth_frame(w, h, seed)
        ex = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
        gk, gd = ex(img)
        rk, rd = oracle.ReferenceExtractor(nf).extract(img)
        assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd)
        ex.close()


@pytest.mark.parametrize("scale,nlevels,w,h", [(3.0, 3, 1200, 900), (2.5, 3, 1000, 800), (2.0, 4, 800, 600), (1.5, 5, 800, 600), (1.1, 8, 640, 480), (1.2, 1, 320, 240)])
def test_other_scale_factors(oracle, gpu_lib, scale, nlevels, w, h):
    """Non-default pyramids: scale 3.0 / 2.5 exercise the generic resize kernel (a pixel's source bytes beyond the reach of the
    word-record selectors), 2.0 and below the word-record kernel at its widest source windows."""
    from ar_orbslam2_b200.extractor import ORBextractor
    img = synth_frame(w, h, 61)
    ex = ORBextractor(800, scale, nlevels, 20, 7, max_width=w, max_height=h)
    ref = oracle.RefExtractor(800, scale, nlevels)
    gk, gd = ex(img)
    rk, rd = ref.extract(img)
    for l in range(nlevels):
        assert np.array_equal(ex.pyramid_level(l, True), ref.pyramid_level(l, True)), l
    assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd)
    ex.close()


def test_opencv24_blur_taps_and_saturation(oracle, gpu_lib):
    """blur_taps=1 selects OpenCV 2.4's rounded-float Gaussian taps {18,34,49,55,49,34,18}.  They sum to 257, so a saturated
    patch reaches 257 before saturate_cast: the image holds white blocks to exercise that clamp."""
    from ar_orbslam2_b200.extractor import ORBextractor
    from ar_orbslam2_b200.synth import synth_frame
    img = synth_frame(640, 480, 12).copy()
    img[100:180, 200:330] = 255
    img[300:420, 40:90] = 255
    img[0:30, 500:640] = 255
    for taps in (1, 0):
        ex = ORBextractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480, blur_taps=taps)
        ref = oracle.RefExtractor(1000, blur_taps=taps)
        gk, gd = ex(img)
        rk, rd = ref.extract(img)
        for l in range(8):
            assert np.array_equal(ex.blurred_level(l), ref.blurred_level(l)), (taps, l)
        if taps == 1:
            assert ref.blurred_level(0).max() == 255
        assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd)
        ex.close()


def test_graph_replay_matches_oracle(oracle, gpu_lib, monkeypatch):
    """Repeated calls at one geometry are captured into a CUDA graph on the second call and replayed afterwards: every
    replay must still equal the oracle (new frame contents flow through the fixed device buffers), a geometry change must
    drop the graph, and ORBGPU_NO_GRAPH=1 must keep the plain launches with the same results and launch counts."""
    from ar_orbslam2_b200.extractor import ORBextractor
    ref = oracle.RefExtractor(1000)
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480)
    per_call = None
    for i, seed in enumerate((31, 32, 33, 34, 35)):
        img = synth_frame(640, 480, seed)
        l0 = ex.launch_count
        gk, gd = ex(img)
        rk, rd = ref.extract(img)
        assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd), f"call {i}"
        per_call = per_call or ex.launch_count - l0
        assert ex.launch_count - l0 == per_call == 34       # 8 x (pyramid, FAST, quadtree, blur) + describe + result export; replays count too
    assert ex.graph_replays == 3                             # call 0 plain, call 1 captured, calls 2-4 replayed
    img = synth_frame(400, 300, 36)                          # other geometry: graph dropped, plain launches
    gk, gd = ex(img)
    rk, rd = ref.extract(img)
    assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd)
    assert ex.graph_replays == 3
    for seed in (37, 38, 39):
        img = synth_frame(400, 300, seed)
        gk, gd = ex(img)
        rk, rd = ref.extract(img)
        assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd)
    assert ex.graph_replays == 5                             # 37 captured, 38 and 39 replayed
    ex.close()
    monkeypatch.setenv("ORBGPU_NO_GRAPH", "1")
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480)
    for seed in (31, 32, 33):
        img = synth_frame(640, 480, seed)
        gk, gd = ex(img)
        rk, rd = ref.extract(img)
        assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd)
    assert ex.graph_replays == 0 and ex.launch_count == 3 * 34
    ex.close()


def test_pitched_host_inputs(oracle, gpu_lib):
    """Host frames that are views into larger buffers: a row pitch that fits the handle's source slot goes up as one linear
    copy, a larger one (and a batch whose frames are not contiguous) through pitched copies; results equal the oracle."""
    import ctypes as C
    from ar_orbslam2_b200 import _lib
    from ar_orbslam2_b200._abi import KEYPOINT_DTYPE, ORB_HOST, ptr
    from ar_orbslam2_b200.extractor import ORBextractor
    w, h = 640, 480
    ref = oracle.RefExtractor(1000)
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=3)
    for extra in (0, 13, 64, 1500):                    # pitch 640 (contiguous), 653 and 704 (linear copy: fit the slot), 2140 (pitched copy)
        buf = np.zeros((h, w + extra), np.uint8)
        img = synth_frame(w, h, 90 + extra)
        off = min(5, extra)
        buf[:, off:off + w] = img
        view = buf[:, off:off + w]
        gk, gd = ex(view)
        rk, rd = ref.extract(img)
        assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd), extra
    # batch of three frames spaced by more than pitch * height (rows of padding between them)
    B = 3
    big = np.zeros((B, h + 9, w), np.uint8)
    imgs = [synth_frame(w, h, 95 + b) for b in range(B)]
    for b in range(B):
        big[b, :h] = imgs[b]
    cap = ex.default_capacity
    kps = np.zeros((B, cap), KEYPOINT_DTYPE)
    desc = np.zeros((B, cap, 32), np.uint8)
    n = np.zeros(B, np.int32)
    L = _lib.lib()
    _lib.check(L.orb_extract_batch(ex._h, ptr(big), B, w, h, w, C.c_size_t((h + 9) * w), ORB_HOST, ptr(kps), ptr(desc), cap, ptr(n),
                                   ORB_HOST), "orb_extract_batch")
    for b in range(B):
        rk, rd = ref.extract(imgs[b])
        assert n[b] == len(rk) and kps[b, : n[b]].tobytes() == rk.tobytes() and np.array_equal(desc[b, : n[b]], rd), b
    ex.close()


def test_two_live_handles_with_different_quotas(oracle, gpu_lib):
    """the quadtree's dynamic shared-memory limit is a per-function attribute.  A smaller handle created AFTER a
    larger one (the reverse of Tracking.cc:455-463) must not lower it under the larger handle's launches."""
    from ar_orbslam2_b200.extractor import ORBextractor
    w, h = 640, 480
    big = ORBextractor(2000, 1.2, 8, 20, 7, max_width=w, max_height=h)
    small = ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h)
    rb, rs = oracle.RefExtractor(2000), oracle.RefExtractor(1000)
    for seed in (1, 2, 3):
        img = synth_frame(w, h, seed)
        for ex, ref in ((big, rb), (small, rs), (big, rb)):
            gk, gd = ex(img)
            rk, rd = ref.extract(img)
            assert gk.tobytes() == rk.tobytes() and np.array_equal(gd, rd)
    big.close(); small.close()


def test_too_many_features_is_a_clear_error(gpu_lib):
    from ar_orbslam2_b200 import _lib
    from ar_orbslam2_b200.extractor import ORBextractor
    with pytest.raises(_lib.OrbGpuError) as e:
        ORBextractor(40000, 1.2, 8, 20, 7, max_width=640, max_height=480)
    assert e.value.code == -1 and "shared memory" in str(e.value)


def test_cvtcolor_chained_in_front_of_extraction(oracle, gpu_lib):
    """a device-mode helper feeding a handle.  The documented way -- pass the handle's stream -- and the thread's
    call stream + orb_call_stream_sync both give the oracle's key points."""
    import ctypes as C
    import torch
    from ar_orbslam2_b200.color import cvtColorGray_device
    from ar_orbslam2_b200.extractor import ORBextractor
    w, h = 640, 480
    g = synth_frame(w, h, 4)
    rgb = np.stack([g, g, g], 2).copy()
    rk, rd = oracle.RefExtractor(1000).extract(g)
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h)
    dev = torch.device("cuda", 0)
    d_rgb = torch.from_numpy(rgb).to(dev)
    cap = ex.default_capacity
    for mode in ("handle_stream", "call_stream"):
        d_grey = torch.zeros((h, w), dtype=torch.uint8, device=dev)
        d_k = torch.zeros((cap, 28), dtype=torch.uint8, device=dev)
        d_d = torch.zeros((cap, 32), dtype=torch.uint8, device=dev)
        d_n = torch.zeros(1, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        if mode == "handle_stream":
            cvtColorGray_device(d_rgb.data_ptr(), w, h, 3 * w, 3 * w * h, 3, True, 1, d_grey.data_ptr(), w, w * h, stream=ex.stream)
        else:
            cvtColorGray_device(d_rgb.data_ptr(), w, h, 3 * w, 3 * w * h, 3, True, 1, d_grey.data_ptr(), w, w * h, stream=None)
            assert gpu_lib.orb_call_stream_sync() == 0
        ex.extract_batch_device(d_grey.data_ptr(), 1, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), cap, d_n.data_ptr())
        ex.sync()
        n = int(d_n.item())
        assert n == len(rk)
        assert d_k[:n].cpu().numpy().tobytes() == rk.tobytes() and np.array_equal(d_d[:n].cpu().numpy(), rd)
    ex.close()


def _edge_frames(w, h, seed):
    rng = np.random.default_rng(seed)
    flat = np.full((h, w), 77, np.uint8)
    noise = rng.integers(0, 256, (h, w), dtype=np.uint8)
    soft = (128 + 20 * rng.standard_normal((h, w))).clip(0, 255).astype(np.uint8)
    ramp = ((np.arange(w)[None, :] * 9 + np.arange(h)[:, None] * 11) % 256).astype(np.uint8)
    half = synth_frame(w, h, seed + 1).copy()
    half[:, w // 2:] = 90   # textured left half, flat right half: non-empty and empty cells side by side
    return [flat, noise, soft, ramp, half]


@pytest.mark.parametrize("mode", ["hybrid", "rows"])
@pytest.mark.parametrize("w,h,nf,nb", [(752, 480, 1000, 12), (1241, 376, 2000, 10), (640, 480, 1500, 14)])
def test_large_batch_fast_paths(oracle, gpu_lib, monkeypatch, w, h, nf, nb, mode):
    """Batches above the latency-mode limit run their large levels through k_fast_rows (one CTA per cell-row segment) and the
    small ones through the per-cell kernel ("hybrid", the default); "rows" forces every level through k_fast_rows.  Same
    candidates per level and same final output as the reference, frame by frame -- flat, dense-noise (list overflow -> per-cell
    routine), soft-noise, both-type ramps and half-empty frames (in-tile retry at minThFAST) included."""
    from ar_orbslam2_b200.extractor import ORBextractor
    if mode == "rows":
        monkeypatch.setenv("ORBGPU_FAST_ROWS_MIN_PX", "0")
    assert nb * w * h > (4 << 20)
    frames = _edge_frames(w, h, 40) + [synth_frame(w, h, 100 + s) for s in range(nb - 5)]
    frames = np.stack(frames)
    ex = ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=nb)
    out = ex.extract_batch(frames)
    ref = oracle.RefExtractor(nf)
    for b in range(nb):
        rk, rd = ref.extract(frames[b])
        for l in range(8):
            a, c = ex.candidates(l, slot=b), ref.candidates(l)
            assert len(a) == len(c), f"frame {b} level {l}: {len(a)} candidates vs {len(c)}"
            assert np.array_equal(a, c), f"frame {b} level {l}: candidate lists differ"
        gk, gd = out[b]
        assert len(gk) == len(rk), (b, len(gk), len(rk))
        assert gk.tobytes() == rk.tobytes(), b
        assert np.array_equal(gd, rd), b
    # a second call on the same handle, frames in another order
    perm = np.roll(np.arange(nb), 3)
    out2 = ex.extract_batch(np.ascontiguousarray(frames[perm]))
    for i, b in enumerate(perm):
        assert out2[i][0].tobytes() == out[b][0].tobytes() and np.array_equal(out2[i][1], out[b][1]), (i, b)
    ex.close()


def test_large_batch_fast_equals_cell_kernel_and_geometry_change(gpu_lib, monkeypatch):
    """k_fast_rows on every level against the per-cell kernel on the same handle geometry, then a smaller geometry on the same
    handle; 25-px-high last cell rows and segments that end inside a word included."""
    from ar_orbslam2_b200.extractor import ORBextractor
    w, h, nb = 960, 540, 10
    frames = np.stack([synth_frame(w, h, 300 + s) for s in range(nb)])
    monkeypatch.setenv("ORBGPU_FAST_ROWS_MIN_PX", "0")
    ex = ORBextractor(1500, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=nb)
    monkeypatch.delenv("ORBGPU_FAST_ROWS_MIN_PX")
    monkeypatch.setenv("ORBGPU_FAST_CELLS", "1")
    ex_cells = ORBextractor(1500, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=nb)
    monkeypatch.delenv("ORBGPU_FAST_CELLS")
    for ww, hh in ((w, h), (900, 500), (933, 517), (w, h)):
        sub = np.ascontiguousarray(frames[:, :hh, :ww])
        a, b = ex.extract_batch(sub), ex_cells.extract_batch(sub)
        for i in range(nb):
            for l in range(8):
                assert np.array_equal(ex.candidates(l, slot=i), ex_cells.candidates(l, slot=i)), (ww, i, l)
            assert a[i][0].tobytes() == b[i][0].tobytes() and np.array_equal(a[i][1], b[i][1]), (ww, i)
    ex.close(); ex_cells.close()


def test_l2_sub_batches_give_the_same_bytes(oracle, gpu_lib, monkeypatch):
    """ORBGPU_CHUNK_MB: a large batch run as sub-batches over four streams (ragged last sub-batch, k_fast_rows with a first-frame
    coordinate in its TMA tile) against the whole batch at once and against the oracle's hash of every frame."""
    from ar_orbslam2_b200.extractor import ORBextractor
    w, h, nb = 752, 480, 27
    frames = np.stack([synth_frame(w, h, 500 + s) for s in range(nb)])
    monkeypatch.setenv("ORBGPU_CHUNK_MB", "16")
    monkeypatch.setenv("ORBGPU_FAST_ROWS_MIN_PX", "0")
    ex = ORBextractor(1200, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=nb)
    monkeypatch.delenv("ORBGPU_CHUNK_MB")
    ex_whole = ORBextractor(1200, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=nb)
    a, b = ex.extract_batch(frames), ex_whole.extract_batch(frames)
    assert ex.launch_count > 2 * ex_whole.launch_count          # it did run as several sub-batches
    for i in range(nb):
        assert a[i][0].tobytes() == b[i][0].tobytes() and np.array_equal(a[i][1], b[i][1]), i
    k, d = oracle.RefExtractor(1200).extract(frames[26])
    assert a[26][0].tobytes() == k.tobytes() and np.array_equal(a[26][1], d)
    ex.close(); ex_whole.close()


def test_l2_sub_batches_with_pinned_export(oracle, gpu_lib, monkeypatch):
    """A sub-batched call whose results are small enough for the pinned export block (12 frames, capacity 788: 567 KB <= 1 MB):
    the sub-batch branch has to fill that block itself, because the host-output tail then copies nothing else.  Run twice with
    different frames so stale contents of the block would show."""
    from ar_orbslam2_b200.extractor import ORBextractor
    w, h, nb = 752, 480, 12
    monkeypatch.setenv("ORBGPU_CHUNK_MB", "1")
    ex = ORBextractor(500, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=nb)
    monkeypatch.delenv("ORBGPU_CHUNK_MB")
    ex_whole = ORBextractor(500, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=nb)
    assert nb * ex.default_capacity * 60 <= 1 << 20 and nb * w * h > 4 << 20
    for first in (700, 740):
        frames = np.stack([synth_frame(w, h, first + s) for s in range(nb)])
        l0 = ex.launch_count, ex_whole.launch_count
        a, b = ex.extract_batch(frames), ex_whole.extract_batch(frames)
        assert ex.launch_count - l0[0] > ex_whole.launch_count - l0[1]      # it did run as sub-batches
        for i in range(nb):
            assert len(a[i][0]) > 400
            assert a[i][0].tobytes() == b[i][0].tobytes() and np.array_equal(a[i][1], b[i][1]), i
        k, d = oracle.RefExtractor(500).extract(frames[nb - 1])
        assert a[nb - 1][0].tobytes() == k.tobytes() and np.array_equal(a[nb - 1][1], d)
    ex.close(); ex_whole.close()
