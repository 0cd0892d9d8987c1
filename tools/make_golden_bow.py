#!/usr/bin/env python3
"""Golden vectors for the DBoW2 transform, produced by the REFERENCE'S OWN DBoW2 (oracle/_ref/libdbow2_ref.so, compiled
unmodified from /root/reference/ORB_SLAM2/Thirdparty/DBoW2 by oracle/Makefile.ref).  Only runs where the reference
checkout exists; the output tests/golden/bow_reference.npz is committed and pins the oracle on the GPU box."""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ar_orbslam2_b200.synth import synth_bow_descriptors, synth_vocabulary, write_vocabulary_text  # noqa: E402
from oracle import orbref  # noqa: E402

# k, L, seed, scoring, weighting, levelsup, n
CASES = [(10, 3, 1, 0, 0, 1, 600), (4, 5, 3, 1, 1, 3, 500), (7, 3, 4, 5, 0, 0, 400), (9, 3, 5, 2, 2, 2, 400), (3, 6, 6, 5, 3, 4, 300)]

assert orbref.ref_dbow_available()
out = {"ncases": np.int32(len(CASES))}
for c, (k, L, seed, scoring, weighting, levelsup, n) in enumerate(CASES):
    voc = synth_vocabulary(k, L, seed=seed)
    d = synth_bow_descriptors(voc, n, seed)
    path = os.path.join(tempfile.mkdtemp(), "voc.txt")
    write_vocabulary_text(path, voc, scoring, weighting, trailing_newline=False)
    R = orbref.ReferenceVocabulary(path)
    out[f"case{c}"] = np.array([k, L, seed, scoring, weighting, levelsup, n], np.int32)
    for name, arr in zip(("bow_word", "bow_value", "fv_node", "fv_off", "fv_feat"), R.transform(d, levelsup)):
        out[f"c{c}_{name}"] = arr
    word, node, w = R.transform_features(d, levelsup)
    out[f"c{c}_word"], out[f"c{c}_node"] = word, node
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "bow_reference.npz"), **out)
print("wrote tests/golden/bow_reference.npz", {k: v.shape for k, v in out.items() if k.startswith("c0_")})
