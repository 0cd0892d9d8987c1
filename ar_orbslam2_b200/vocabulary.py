"""Host-side mirror of ORB_SLAM2::ORBVocabulary (ref: ORB_SLAM2/include/ORBVocabulary.h:29 =
DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h) over liborbgpu:
the tree lives in HBM, `transform` is Frame::ComputeBoW / KeyFrame::ComputeBoW (src/Frame.cc:400-407,
src/KeyFrame.cc:62-70).  No CPU fallback."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._abi import ORB_DEVICE, ORB_HOST, ptr

# DBoW2::WeightingType / ScoringType (BowVector.h:36-55)
TF_IDF, TF, IDF, BINARY = 0, 1, 2, 3
L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT = range(6)


class OrbVocabDesc(C.Structure):
    _fields_ = [("k", C.c_int32), ("L", C.c_int32), ("scoring", C.c_int32), ("weighting", C.c_int32), ("n_nodes", C.c_int32),
                ("parent", C.c_void_p), ("descriptors", C.c_void_p), ("weight", C.c_void_p), ("word_id", C.c_void_p)]


class OrbBowOut(C.Structure):
    _fields_ = [("capacity", C.c_int32), ("bow_word", C.c_void_p), ("bow_value", C.c_void_p), ("n_bow", C.c_void_p),
                ("fv_node", C.c_void_p), ("fv_off", C.c_void_p), ("fv_feat", C.c_void_p), ("n_fv", C.c_void_p)]


def _setup(L):
    L.orb_vocab_destroy.argtypes = [C.c_void_p]
    L.orb_vocab_destroy.restype = None
    L.orb_vocab_stream.argtypes = [C.c_void_p]
    L.orb_vocab_stream.restype = C.c_void_p
    L.orb_vocab_launch_count.argtypes = [C.c_void_p]
    L.orb_vocab_launch_count.restype = C.c_int64
    return L


class ORBVocabulary:
    """`ORBVocabulary()` then `loadFromTextFile(path)` like src/System.cc:67-75, or `ORBVocabulary.from_arrays(...)`."""

    def __init__(self, device: int = -1):
        self._L = _setup(_lib.lib())
        self.device = device
        self.h = None
        _lib.register(self)

    # -- construction -------------------------------------------------------------------------------------------------
    def loadFromTextFile(self, path: str) -> bool:
        """TemplatedVocabulary::loadFromTextFile (TemplatedVocabulary.h:1338-1425)."""
        self.close()
        h = C.c_void_p()
        rc = self._L.orb_vocab_load_text(str(path).encode(), self.device, C.byref(h))
        if rc == -1:     # ORB_ERR_INVALID: the reference returns false on a file that is not a vocabulary
            return False
        _lib.check(rc, "orb_vocab_load_text")
        self.h = h
        return True

    @classmethod
    def from_arrays(cls, k, L, parent, descriptors, weight, word_id, scoring=L1_NORM, weighting=TF_IDF, device=-1):
        self = cls(device)
        parent = np.ascontiguousarray(parent, np.int32)
        descriptors = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        weight = np.ascontiguousarray(weight, np.float64)
        word_id = np.ascontiguousarray(word_id, np.int32)
        n = parent.shape[0]
        assert descriptors.shape[0] == n and weight.shape[0] == n and word_id.shape[0] == n
        d = OrbVocabDesc(k, L, scoring, weighting, n, parent.ctypes.data, descriptors.ctypes.data, weight.ctypes.data, word_id.ctypes.data)
        h = C.c_void_p()
        _lib.check(self._L.orb_vocab_create(C.byref(d), device, C.byref(h)), "orb_vocab_create")
        self.h = h
        return self

    def close(self):
        if getattr(self, "h", None):
            self._L.orb_vocab_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- introspection ------------------------------------------------------------------------------------------------
    def info(self):
        a = np.zeros(6, np.int32)
        _lib.check(self._L.orb_vocab_info(self.h, ptr(a)), "orb_vocab_info")
        return dict(k=int(a[0]), L=int(a[1]), scoring=int(a[2]), weighting=int(a[3]), nodes=int(a[4]), words=int(a[5]))

    def size(self) -> int:
        return self.info()["words"]

    def empty(self) -> bool:
        return self.size() == 0

    def nodes(self):
        n = self.info()["nodes"]
        parent, word = np.zeros(n, np.int32), np.zeros(n, np.int32)
        desc, weight = np.zeros((n, 32), np.uint8), np.zeros(n, np.float64)
        _lib.check(self._L.orb_vocab_nodes(self.h, ptr(parent), ptr(desc), ptr(weight), ptr(word)), "orb_vocab_nodes")
        return parent, desc, weight, word

    @property
    def launch_count(self) -> int:
        return int(self._L.orb_vocab_launch_count(self.h))

    @property
    def stream(self) -> int:
        return int(self._L.orb_vocab_stream(self.h) or 0)

    # -- transform ----------------------------------------------------------------------------------------------------
    def transform_features(self, descriptors: np.ndarray, levelsup: int = 0):
        """Per-feature transform (TemplatedVocabulary.h:1217-1255): (word id, node id `levelsup` levels up, weight)."""
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        n = d.shape[0]
        word, node, w = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float64)
        _lib.check(self._L.orb_vocab_transform_features(self.h, ptr(d), n, ORB_HOST, levelsup, ptr(word), ptr(node), ptr(w)),
                   "orb_vocab_transform_features")
        return word, node, w

    def transform(self, descriptors: np.ndarray, levelsup: int = 4):
        """transform(features, BowVector&, FeatureVector&, levelsup) (TemplatedVocabulary.h:1126-1193) for one frame.
        Returns (bow_word[int32], bow_value[float64], fv_node[int32], fv_off[int32], fv_feat[int32]): the two std::maps in
        iteration order, the FeatureVector as CSR (the form SearchByBoW / SearchForTriangulation take)."""
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        r = self.transform_batch(d[None] if d.shape[0] else np.zeros((1, 1, 32), np.uint8), np.array([d.shape[0]], np.int32), levelsup)
        return r[0]

    def transform_batch(self, descriptors: np.ndarray, counts: np.ndarray, levelsup: int = 4):
        """B frames in one call: descriptors [B, stride, 32], counts [B].  Returns a list of per-frame tuples."""
        d = np.ascontiguousarray(descriptors, np.uint8)
        B, stride = d.shape[0], d.shape[1]
        n = np.ascontiguousarray(counts, np.int32)
        cap = max(1, int(n.max()) if B else 1)
        bw, bv, nb = np.zeros((B, cap), np.int32), np.zeros((B, cap), np.float64), np.zeros(B, np.int32)
        fn, fo, ff, nf = np.zeros((B, cap), np.int32), np.zeros((B, cap + 1), np.int32), np.zeros((B, cap), np.int32), np.zeros(B, np.int32)
        out = OrbBowOut(cap, bw.ctypes.data, bv.ctypes.data, nb.ctypes.data, fn.ctypes.data, fo.ctypes.data, ff.ctypes.data, nf.ctypes.data)
        _lib.check(self._L.orb_vocab_transform(self.h, ptr(d), ptr(n), B, stride, ORB_HOST, levelsup, C.byref(out), ORB_HOST),
                   "orb_vocab_transform")
        res = []
        for b in range(B):
            g = int(nf[b])
            res.append((bw[b, : nb[b]].copy(), bv[b, : nb[b]].copy(), fn[b, :g].copy(), fo[b, : g + 1].copy(), ff[b, : fo[b, g]].copy()))
        return res

    def transform_device(self, d_desc: int, d_counts: int, batch: int, stride: int, levelsup: int, capacity: int,
                         d_bow_word: int, d_bow_value: int, d_n_bow: int, d_fv_node: int, d_fv_off: int, d_fv_feat: int, d_n_fv: int):
        """Device-resident form, asynchronous on the vocabulary's stream (raw device addresses)."""
        out = OrbBowOut(capacity, d_bow_word, d_bow_value, d_n_bow, d_fv_node, d_fv_off, d_fv_feat, d_n_fv)
        _lib.check(self._L.orb_vocab_transform(self.h, C.c_void_p(d_desc), C.c_void_p(d_counts), batch, stride, ORB_DEVICE, levelsup,
                                               C.byref(out), ORB_DEVICE), "orb_vocab_transform")

    def sync(self):
        _lib.check(self._L.orb_vocab_sync(self.h), "orb_vocab_sync")
