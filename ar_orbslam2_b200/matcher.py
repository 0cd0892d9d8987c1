"""Host-side mirror of ORB_SLAM2::ORBmatcher (ref: ORB_SLAM2/include/ORBmatcher.h:34-114) over liborbgpu:
the Hamming distance, the brute-force top-2 scan and the keyframe-range-sharded descriptor database
(config C5).  The windowed tracking matchers live in tracking.py."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._abi import ORB_DEVICE, ORB_HOST, ptr

TH_HIGH = 100   # ref ORBmatcher.cc:37
TH_LOW = 50     # ref ORBmatcher.cc:38
HISTO_LENGTH = 30


def DescriptorDistance(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """ORBmatcher::DescriptorDistance (ref ORBmatcher.cc:1650-1666) for n descriptor pairs, on the GPU."""
    a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
    b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
    assert a.shape == b.shape
    out = np.zeros(a.shape[0], np.int32)
    _lib.check(_lib.lib().orb_descriptor_distance(ptr(a), ptr(b), a.shape[0], ptr(out), ORB_HOST), "orb_descriptor_distance")
    return out


def hamming_top2(queries: np.ndarray, db: np.ndarray, idx_base: int = 0):
    """Best and second-best database row per query (strict '<', first row wins ties).  Returns
    (best_dist, best_idx, second_dist, second_idx); 256 / -1 where absent."""
    q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 32)
    d = np.ascontiguousarray(db, np.uint8).reshape(-1, 32)
    nq = q.shape[0]
    d1, d2 = np.zeros(nq, np.int32), np.zeros(nq, np.int32)
    i1, i2 = np.zeros(nq, np.int64), np.zeros(nq, np.int64)
    _lib.check(_lib.lib().orb_hamming_top2(ptr(q), nq, ptr(d), C.c_int64(d.shape[0]), C.c_int64(idx_base), ptr(d1), ptr(i1),
                                           ptr(d2), ptr(i2), ORB_HOST), "orb_hamming_top2")
    return d1, i1, d2, i2


TOP2_STATIC_MAX_QUERIES = 8192   # ORB_TOP2_STATIC_MAX_QUERIES of include/orbgpu.h


def hamming_top2_static_device(d_queries, nq, d_db, ndb, idx_base, o_bd, o_bi, o_sd, o_si, stream=0):
    """orb_hamming_top2_static: the brute-force top-2 scan for callers that must not allocate device memory while they
    run (a tracking loop with memory fixed at start-up, a host whose own pool owns the GPU heap).  Every argument is a raw
    device address; the chunk partials go to a fixed 6 MiB array inside liborbgpu instead of a cudaMalloc'ed block, so at
    most TOP2_STATIC_MAX_QUERIES queries per call, and calls that can overlap on one GPU must be ordered by the caller.
    Enqueues on `stream` (a cudaStream_t handle) and returns without waiting."""
    vp = C.c_void_p
    _lib.check(_lib.lib().orb_hamming_top2_static(vp(d_queries), int(nq), vp(d_db), C.c_int64(ndb), C.c_int64(idx_base), vp(o_bd),
                                                  vp(o_bi), vp(o_sd), vp(o_si), vp(stream)), "orb_hamming_top2_static")


def top2_merge(best_dist, best_idx, second_dist, second_idx):
    """Merge [nshards, nq] partial lists (shards in database order) into the single-scan result."""
    bd = np.ascontiguousarray(best_dist, np.int32)
    bi = np.ascontiguousarray(best_idx, np.int64)
    sd = np.ascontiguousarray(second_dist, np.int32)
    si = np.ascontiguousarray(second_idx, np.int64)
    ns, nq = bd.shape
    o = (np.zeros(nq, np.int32), np.zeros(nq, np.int64), np.zeros(nq, np.int32), np.zeros(nq, np.int64))
    _lib.check(_lib.lib().orb_top2_merge(ns, nq, ptr(bd), ptr(bi), ptr(sd), ptr(si), ptr(o[0]), ptr(o[1]), ptr(o[2]), ptr(o[3]),
                                         ORB_HOST), "orb_top2_merge")
    return o


def top2_merge_device(nshards, nq, d_bd, d_bi, d_sd, d_si, o_bd, o_bi, o_sd, o_si):
    vp = C.c_void_p
    _lib.check(_lib.lib().orb_top2_merge(nshards, nq, vp(d_bd), vp(d_bi), vp(d_sd), vp(d_si), vp(o_bd), vp(o_bi), vp(o_sd),
                                         vp(o_si), ORB_DEVICE), "orb_top2_merge")


def measure_popc_rate() -> float:
    """Measured POPC throughput of the device in Gpopc32/s (matching roofline denominator)."""
    v = C.c_double()
    _lib.check(_lib.lib().orb_measure_popc_rate(C.byref(v)), "orb_measure_popc_rate")
    return v.value


class DescriptorDB:
    """Shard [row_begin,row_end) of a global descriptor database resident in HBM (keyframe-range sharding,
    replaces the BoW inverted index of ref KeyFrameDatabase.cc:185 with a brute-force scan)."""

    def __init__(self, row_begin: int, row_end: int, device: int = -1):
        self._L = _lib.lib()
        self._L.orb_db_device_ptr.restype = C.c_void_p
        self._L.orb_db_stream.restype = C.c_void_p
        self._L.orb_db_launch_count.restype = C.c_int64
        self._h = C.c_void_p()
        _lib.check(self._L.orb_db_create(C.c_int64(row_begin), C.c_int64(row_end), device, C.byref(self._h)), "orb_db_create")
        self.row_begin, self.row_end = row_begin, row_end
        _lib.register(self)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orb_db_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, first_row: int, rows: np.ndarray):
        rows = np.ascontiguousarray(rows, np.uint8).reshape(-1, 32)
        _lib.check(self._L.orb_db_upload(self._h, C.c_int64(first_row), ptr(rows), C.c_int64(rows.shape[0]), ORB_HOST), "orb_db_upload")

    def fill_synthetic(self, seed: int):
        _lib.check(self._L.orb_db_fill_synthetic(self._h, C.c_uint64(seed)), "orb_db_fill_synthetic")

    def row(self, global_row: int) -> np.ndarray:
        out = np.zeros(32, np.uint8)
        _lib.check(self._L.orb_db_row(self._h, C.c_int64(global_row), ptr(out)), "orb_db_row")
        return out

    def rows(self, idx) -> np.ndarray:
        return np.stack([self.row(int(i)) for i in idx]) if len(idx) else np.zeros((0, 32), np.uint8)

    @property
    def stream(self) -> int:
        return self._L.orb_db_stream(self._h)

    @property
    def launch_count(self) -> int:
        return self._L.orb_db_launch_count(self._h)

    def query(self, queries: np.ndarray):
        q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 32)
        nq = q.shape[0]
        d1, d2 = np.zeros(nq, np.int32), np.zeros(nq, np.int32)
        i1, i2 = np.zeros(nq, np.int64), np.zeros(nq, np.int64)
        _lib.check(self._L.orb_db_query(self._h, ptr(q), nq, ptr(d1), ptr(i1), ptr(d2), ptr(i2), ORB_HOST), "orb_db_query")
        return d1, i1, d2, i2

    def query_device(self, d_q: int, nq: int, d_d1: int, d_i1: int, d_d2: int, d_i2: int):
        """Asynchronous on the shard's stream; raw device pointers."""
        vp = C.c_void_p
        _lib.check(self._L.orb_db_query(self._h, vp(d_q), nq, vp(d_d1), vp(d_i1), vp(d_d2), vp(d_i2), ORB_DEVICE), "orb_db_query")

    def query_keyframes(self, queries: np.ndarray, rows_per_keyframe: int, th_low: int = 50, ratio: float = 0.75):
        """Shard top-2 plus, per keyframe of the shard, the number of queries it matches (best <= th_low and
        best < ratio * second inside the keyframe; ref ORBmatcher.cc:216-231).  Host arrays in, host arrays out."""
        q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 32)
        nq = q.shape[0]
        d1, d2 = np.zeros(nq, np.int32), np.zeros(nq, np.int32)
        i1, i2 = np.zeros(nq, np.int64), np.zeros(nq, np.int64)
        counts = np.zeros((self.row_end - self.row_begin) // rows_per_keyframe, np.int32)
        _lib.check(self._L.orb_db_query_keyframes(self._h, ptr(q), nq, rows_per_keyframe, th_low, C.c_float(ratio), ptr(d1), ptr(i1),
                                                  ptr(d2), ptr(i2), ptr(counts), ORB_HOST), "orb_db_query_keyframes")
        return d1, i1, d2, i2, counts

    def query_keyframes_device(self, d_q: int, nq: int, rows_per_keyframe: int, th_low: int, ratio: float, d_d1: int, d_i1: int,
                               d_d2: int, d_i2: int, d_counts: int):
        """Asynchronous on the shard's stream; raw device pointers."""
        vp = C.c_void_p
        _lib.check(self._L.orb_db_query_keyframes(self._h, vp(d_q), nq, rows_per_keyframe, th_low, C.c_float(ratio), vp(d_d1), vp(d_i1),
                                                  vp(d_d2), vp(d_i2), vp(d_counts), ORB_DEVICE), "orb_db_query_keyframes")

    def merge_top2_device(self, nshards: int, nq: int, d_bd: int, d_bi: int, d_sd: int, d_si: int, o_bd: int, o_bi: int, o_sd: int, o_si: int):
        """[nshards, nq] partial lists (device pointers, shard order) -> single-scan result, on the shard's stream."""
        vp = C.c_void_p
        _lib.check(self._L.orb_db_merge_top2(self._h, nshards, nq, vp(d_bd), vp(d_bi), vp(d_sd), vp(d_si), vp(o_bd), vp(o_bi), vp(o_sd),
                                             vp(o_si)), "orb_db_merge_top2")


def ComputeDistinctiveDescriptors(descriptors: np.ndarray, offsets: np.ndarray):
    """MapPoint::ComputeDistinctiveDescriptors (ref MapPoint.cc:224-285) for many map points at once, on the GPU: map point
    p is observed through descriptor rows [offsets[p], offsets[p+1]) (non-bad key frames in mObservations order).
    Returns (best_row, best_median): the index inside each group of the descriptor with the least median distance to
    the others (-1 for an empty group)."""
    d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
    off = np.ascontiguousarray(offsets, np.int32)
    n = off.shape[0] - 1
    row, med = np.zeros(n, np.int32), np.zeros(n, np.int32)
    _lib.check(_lib.lib().orb_distinctive_descriptors(ptr(d), ptr(off), n, ORB_HOST, ptr(row), ptr(med)), "orb_distinctive_descriptors")
    return row, med


def searchMatches(descp1: np.ndarray, descp2: np.ndarray, ratio: float = 0.8, max_dist: int = 50):
    """Marker::searchMatches (ref Marker.cc:314-349) / naive_nn_search2 (AR-1.3/src/ORBMatcher.cpp:70-102) on the GPU: descp1 =
    train (marker) descriptors, descp2 = query (frame) descriptors.  Returns the DMatch list as arrays
    (queryIdx, trainIdx, distance) in the reference's push_back order."""
    t = np.ascontiguousarray(descp1, np.uint8).reshape(-1, 32)
    q = np.ascontiguousarray(descp2, np.uint8).reshape(-1, 32)
    idx, dist = np.zeros(len(q), np.int32), np.zeros(len(q), np.int32)
    _lib.check(_lib.lib().orb_search_matches_ratio(ptr(t), len(t), ptr(q), len(q), C.c_double(ratio), max_dist, ORB_HOST, ptr(idx),
                                                   ptr(dist)), "orb_search_matches_ratio")
    keep = np.nonzero(idx >= 0)[0].astype(np.int32)
    return keep, idx[keep], dist[keep].astype(np.float32)


def hamming_minmax(descp1: np.ndarray, descp2: np.ndarray):
    """Smallest / largest Hamming distance over all (query = descp2, train = descp1) pairs: naive_nn_search2's minD / maxD
    bookkeeping (ref AR-1.3/src/ORBMatcher.cpp:92-93).  (INT_MAX, -1) without a pair."""
    t = np.ascontiguousarray(descp1, np.uint8).reshape(-1, 32)
    q = np.ascontiguousarray(descp2, np.uint8).reshape(-1, 32)
    mm = np.zeros(2, np.int32)
    _lib.check(_lib.lib().orb_hamming_minmax(ptr(t), len(t), ptr(q), len(q), ORB_HOST, ptr(mm)), "orb_hamming_minmax")
    return int(mm[0]), int(mm[1])
