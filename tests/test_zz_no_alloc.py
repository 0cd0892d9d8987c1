"""orb_hamming_top2_static -- the brute-force top-2 scan for callers that cannot allocate GPU memory -- against
orb_hamming_top2 through the C ABI.

Some users of liborbgpu may not call cudaMalloc while they run: a relocalisation query inside a real-time tracking loop whose
memory was sized at start-up, a host that has given the whole device heap to its own pool, a process so close to its memory
limit that an allocation in mid-frame could fail.  For them the library keeps the scan's chunk partials in a fixed 6 MiB device
array instead of a block it grows on demand.  That array is shared by every call on a GPU and is never cleared, so the test makes
two calls in a row with different shapes -- the second smaller, so stale partials of the first are still in the array -- and asks
for the same bytes as the allocating entry point both times.  Integer results: the tolerance is zero.

This file is named to be collected last: `pytest -x` stops at the first failure, and a failure here must not hide the others."""
import ctypes as C

import numpy as np
import pytest

from ar_orbslam2_b200.synth import synth_descriptors

pytestmark = pytest.mark.gpu


def _existing_entry_point(lib, d_q, nq, d_db, ndb, idx_base, torch):
    """orb_hamming_top2 on the same device buffers (it allocates its scratch; it waits before returning)."""
    from ar_orbslam2_b200 import _lib
    from ar_orbslam2_b200._abi import ORB_DEVICE
    dev = d_q.device
    out = (torch.empty(nq, dtype=torch.int32, device=dev), torch.empty(nq, dtype=torch.int64, device=dev),
           torch.empty(nq, dtype=torch.int32, device=dev), torch.empty(nq, dtype=torch.int64, device=dev))
    vp = C.c_void_p
    torch.cuda.synchronize()
    _lib.check(lib.orb_hamming_top2(vp(d_q.data_ptr()), nq, vp(d_db.data_ptr() if ndb else 0), C.c_int64(ndb), C.c_int64(idx_base),
                                    vp(out[0].data_ptr()), vp(out[1].data_ptr()), vp(out[2].data_ptr()), vp(out[3].data_ptr()),
                                    ORB_DEVICE), "orb_hamming_top2")
    return [t.cpu().numpy() for t in out]


def _no_alloc_entry_point(d_q, nq, d_db, ndb, idx_base, stream, torch):
    from ar_orbslam2_b200.matcher import hamming_top2_static_device
    dev = d_q.device
    # outputs pre-filled with a pattern no result can hold, so an entry the kernels did not write shows up
    out = (torch.full((nq,), -7, dtype=torch.int32, device=dev), torch.full((nq,), -7, dtype=torch.int64, device=dev),
           torch.full((nq,), -7, dtype=torch.int32, device=dev), torch.full((nq,), -7, dtype=torch.int64, device=dev))
    torch.cuda.synchronize()
    hamming_top2_static_device(d_q.data_ptr(), nq, d_db.data_ptr() if ndb else 0, ndb, idx_base, out[0].data_ptr(), out[1].data_ptr(),
                               out[2].data_ptr(), out[3].data_ptr(), stream.cuda_stream)
    stream.synchronize()
    return [t.cpu().numpy() for t in out]


def test_no_alloc_top2_equals_allocating_top2_twice_in_a_row(gpu_lib):
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.Stream(device=dev)
    # call 1: 2000 queries (the chunk count is held down to fit the fixed array), half of them near-copies of database rows;
    # call 2: fewer queries AND fewer rows, so most of the array still holds call 1's partials; call 3: an empty database
    shapes = [(2000, 60_000, 11, 1_000_000), (300, 777, 12, 0), (5, 0, 13, 42)]
    for nq, ndb, seed, idx_base in shapes:
        db = synth_descriptors(max(ndb, 1), seed)[:ndb]
        q = synth_descriptors(nq, seed + 100)
        if ndb:
            rng = np.random.default_rng(seed)
            pick = rng.integers(0, ndb, nq // 2)
            q[: nq // 2] = db[pick]
            q[: nq // 2, rng.integers(0, 32)] ^= np.uint8(1 << int(rng.integers(0, 8)))
            db[ndb // 3] = db[ndb // 7]                      # duplicate rows: the lower index has to win the tie
        d_q = torch.from_numpy(q).to(dev)
        d_db = torch.from_numpy(db if ndb else np.zeros((1, 32), np.uint8)).to(dev)
        want = _existing_entry_point(gpu_lib, d_q, nq, d_db, ndb, idx_base, torch)
        for attempt in range(2):                             # the same call twice: its own leftovers are in the array too
            got = _no_alloc_entry_point(d_q, nq, d_db, ndb, idx_base, stream, torch)
            for g, w, name in zip(got, want, ("best_dist", "best_idx", "second_dist", "second_idx")):
                assert g.dtype == w.dtype and np.array_equal(g, w), f"{name} differs (nq={nq}, ndb={ndb}, attempt {attempt})"
        if ndb == 0:
            assert (want[0] == 256).all() and (want[1] == -1).all()


def test_no_alloc_top2_rejects_what_does_not_fit(gpu_lib):
    import torch
    from ar_orbslam2_b200 import _lib
    from ar_orbslam2_b200.matcher import TOP2_STATIC_MAX_QUERIES, hamming_top2_static_device
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    nq = TOP2_STATIC_MAX_QUERIES + 1
    d_q = torch.zeros((nq, 32), dtype=torch.uint8, device=dev)
    d_db = torch.zeros((64, 32), dtype=torch.uint8, device=dev)
    o32 = torch.full((nq,), -7, dtype=torch.int32, device=dev)
    o64 = torch.full((nq,), -7, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    with pytest.raises(_lib.OrbGpuError) as e:
        hamming_top2_static_device(d_q.data_ptr(), nq, d_db.data_ptr(), 64, 0, o32.data_ptr(), o64.data_ptr(), o32.data_ptr(), o64.data_ptr())
    assert e.value.code == -4 and "8192" in str(e.value)     # ORB_ERR_CAPACITY, and the message names the limit
    torch.cuda.synchronize()
    assert (o32 == -7).all() and (o64 == -7).all()           # nothing was launched
    # the largest size that does fit runs, and agrees with the allocating entry point
    nq = TOP2_STATIC_MAX_QUERIES
    q = synth_descriptors(nq, 5)
    db = synth_descriptors(5000, 6)
    d_q, d_db = torch.from_numpy(q).to(dev), torch.from_numpy(db).to(dev)
    stream = torch.cuda.Stream(device=dev)
    want = _existing_entry_point(gpu_lib, d_q, nq, d_db, 5000, 0, torch)
    got = _no_alloc_entry_point(d_q, nq, d_db, 5000, 0, stream, torch)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
