"""Generates tests/golden/vocab_ref.npz from the REFERENCE's own DBoW2 (oracle/_ref/libdbow2_ref.so, built by
`make -C oracle ref` from /root/reference/Thirdparty/DBoW2): small synthetic vocabularies written in the ORBvoc.txt
format, loaded with the reference's loadFromTextFile and transformed with its transform(features, v, fv, levelsup).
Run in the build container (needs /root/reference); the .npz travels to the GPU box."""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from orb_slam_b200.synth import synth_vocabulary, vocabulary_text  # noqa: E402
from tests.oracle_lib import RefVocabulary  # noqa: E402
from tests.scenes import unbalanced_vocabulary, vocabulary_queries  # noqa: E402

CASES = [  # name, builder, k, L, scoring, weighting, levelsup, nqueries
    ("balanced_k8_L3_tfidf_l1", lambda: synth_vocabulary(8, 3, seed=11, stop_fraction=0.05), 8, 3, 0, 0, 2, 400),
    ("ragged_k7_L4_tfidf_l1", lambda: unbalanced_vocabulary(7, 4, seed=3, min_leaf_level=2), 7, 4, 0, 0, 2, 400),
    ("ragged_k7_L4_tf_l2", lambda: unbalanced_vocabulary(7, 4, seed=4, min_leaf_level=3), 7, 4, 1, 1, 1, 300),
    ("ragged_k5_L5_idf_dot", lambda: unbalanced_vocabulary(5, 5, seed=5), 5, 5, 5, 2, 4, 300),
    ("ragged_k5_L5_binary_l1", lambda: unbalanced_vocabulary(5, 5, seed=6), 5, 5, 0, 3, 7, 300),
    ("ragged_k6_L3_tf_dot", lambda: unbalanced_vocabulary(6, 3, seed=7, min_leaf_level=3), 6, 3, 5, 1, 0, 300),
]


def main():
    out = {}
    for name, build, k, L, scoring, weighting, levelsup, nq in CASES:
        parent, leaf, desc, weight = build()
        q = vocabulary_queries(desc, leaf, nq, seed=len(name))
        with tempfile.NamedTemporaryFile("w", suffix=".txt", delete=False) as f:
            f.write(vocabulary_text(k, L, parent, leaf, desc, weight, scoring, weighting))
            path = f.name
        ref = RefVocabulary(path)
        wid, wv, fv = ref.transform(q, levelsup)
        os.unlink(path)
        out[name + "/meta"] = np.array([k, L, scoring, weighting, levelsup], np.int32)
        out[name + "/parent"] = parent; out[name + "/leaf"] = leaf; out[name + "/desc"] = desc; out[name + "/weight"] = weight
        out[name + "/queries"] = q; out[name + "/word_ids"] = wid; out[name + "/word_vals"] = wv
        ids = np.array(sorted(fv), np.uint32)
        out[name + "/node_ids"] = ids
        out[name + "/node_start"] = np.cumsum([0] + [len(fv[int(i)]) for i in ids]).astype(np.int32)
        out[name + "/items"] = np.array([x for i in ids for x in fv[int(i)]], np.uint32)
        print(name, "nodes", len(parent), "words", ref.size(), "bow", len(wid), "fv nodes", len(ids))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "vocab_ref.npz"), **out)


if __name__ == "__main__":
    main()
