"""GPU parity of the vocabulary transform (orbm_vocab_*: Frame::ComputeBoW, src/Frame.cc:521-528 ->
Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1126-1259) and of orbm_distinctive_descriptors
(MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:241-306): against the golden vectors of the reference's own
DBoW2 build, and against the CPU oracle at the full ORBvoc shape (k = 10, L = 6).  Word ids, node ids, feature lists
and the BowVector doubles are compared bit for bit."""
import os

import numpy as np
import pytest

from orb_slam_b200.synth import synth_vocabulary, vocabulary_text
from tests.oracle_lib import OracleVocabulary, oracle_distinctive_descriptors
from tests.scenes import observation_lists, unbalanced_vocabulary, vocabulary_queries

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "vocab_ref.npz")


@pytest.fixture(scope="module")
def ctx():
    from orb_slam_b200.matcher import MatcherContext
    c = MatcherContext(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def orbvoc_shape(ctx):
    from orb_slam_b200.vocabulary import ORBVocabulary
    arrays = synth_vocabulary(10, 6)
    g = ORBVocabulary.from_arrays(ctx, 10, 6, *arrays)
    o = OracleVocabulary.from_arrays(10, 6, *arrays)
    yield arrays, g, o
    g.close()


def same_transform(a, b):
    assert np.array_equal(a[0], b[0])
    assert np.array_equal(a[1], b[1])            # float64, bit for bit
    fa = a[2].as_dict() if hasattr(a[2], "as_dict") else a[2]
    fb = b[2].as_dict() if hasattr(b[2], "as_dict") else b[2]
    assert fa == fb


def golden_names():
    z = np.load(GOLD)
    return sorted({k.split("/")[0] for k in z.files})


@pytest.mark.parametrize("name", golden_names())
def test_transform_matches_reference_dbow2_golden(ctx, name):
    from orb_slam_b200.vocabulary import ORBVocabulary
    z = np.load(GOLD)
    k, L, scoring, weighting, levelsup = (int(v) for v in z[name + "/meta"])
    voc = ORBVocabulary.from_arrays(ctx, k, L, z[name + "/parent"], z[name + "/leaf"], z[name + "/desc"], z[name + "/weight"],
                                    scoring, weighting)
    wid, wv, fv = voc.transform(z[name + "/queries"], levelsup)
    ids, st, it = z[name + "/node_ids"], z[name + "/node_start"], z[name + "/items"]
    same_transform((wid, wv, fv), (z[name + "/word_ids"], z[name + "/word_vals"],
                                   {int(ids[j]): it[st[j]:st[j + 1]].tolist() for j in range(len(ids))}))
    voc.close()


@pytest.mark.parametrize("n,levelsup", [(2000, 4), (1, 4), (37, 0), (5000, 6), (3000, 2), (2000, 9)])
def test_transform_orbvoc_shape(orbvoc_shape, n, levelsup):
    (parent, leaf, desc, weight), g, o = orbvoc_shape
    assert g.size() == 10 ** 6 and g.nnodes == 1111111
    q = vocabulary_queries(desc, leaf, n, seed=n + levelsup)
    a, b = g.transform(q, levelsup), o.transform(q, levelsup)
    same_transform(a, b)
    assert abs(a[1].sum() - 1.0) < 1e-9          # L1-normalised
    w, nid, wt = g.descend(q, levelsup)
    ow, onid, owt = o.descend(q, levelsup)
    assert np.array_equal(w, ow) and np.array_equal(nid, onid) and np.array_equal(wt, owt)
    if levelsup == 4:
        assert len(a[2].node_ids) <= 100 and (wt == 0).any() == (len(a[2].items) < n)


def test_transform_on_extracted_descriptors(orbvoc_shape, ctx):
    """Frame::ComputeBoW on what the extractor produces for a KITTI-sized frame."""
    from orb_slam_b200 import ORBextractor
    from orb_slam_b200.synth import synth_frame
    _, g, o = orbvoc_shape
    ex = ORBextractor(2000, 1.2, 8, 20, 7, device=0)
    _, d = ex(synth_frame(1241, 376, 3))
    ex.close()
    same_transform(g.transform(d, 4), o.transform(d, 4))


@pytest.mark.parametrize("weighting,scoring", [(0, 0), (1, 1), (2, 5), (3, 0), (1, 5), (0, 3)])
def test_weighting_and_scoring_types(ctx, weighting, scoring):
    from orb_slam_b200.vocabulary import ORBVocabulary
    arrays = unbalanced_vocabulary(9, 5, seed=40 + weighting, min_leaf_level=2)
    g = ORBVocabulary.from_arrays(ctx, 9, 5, *arrays, scoring=scoring, weighting=weighting)
    o = OracleVocabulary.from_arrays(9, 5, *arrays, scoring=scoring, weighting=weighting)
    q = vocabulary_queries(arrays[2], arrays[1], 1200, seed=weighting)
    for levelsup in (3, 5):
        same_transform(g.transform(q, levelsup), o.transform(q, levelsup))
    g.close()


def test_text_loader(ctx, tmp_path):
    from orb_slam_b200.vocabulary import ORBVocabulary
    arrays = unbalanced_vocabulary(8, 4, seed=77, min_leaf_level=2)
    for trailing in (False, True):
        path = tmp_path / ("voc%d.txt" % trailing)
        path.write_text(vocabulary_text(8, 4, *arrays, scoring=0, weighting=0, trailing_newline=trailing))
        g = ORBVocabulary.loadFromTextFile(ctx, path)
        o = OracleVocabulary.loadFromTextFile(path)
        assert (g.k, g.L, g.nnodes, g.nwords) == (8, 4, len(arrays[0]), int(arrays[1].sum())) and o.size() == g.nwords
        q = vocabulary_queries(arrays[2], arrays[1], 700, seed=3)
        same_transform(g.transform(q, 2), o.transform(q, 2))
        g.close()


def test_edge_cases(ctx):
    from orb_slam_b200 import FrontendError
    from orb_slam_b200.vocabulary import ORBVocabulary
    arrays = unbalanced_vocabulary(4, 3, seed=5)
    g = ORBVocabulary.from_arrays(ctx, 4, 3, *arrays)
    wid, wv, fv = g.transform(np.zeros((0, 32), np.uint8), 1)                 # no features
    assert len(wid) == 0 and len(wv) == 0 and len(fv.node_ids) == 0
    g.close()
    empty = ORBVocabulary.from_arrays(ctx, 10, 6, [0], [0], np.zeros((1, 32), np.uint8), [0.0])    # empty(): nothing returned
    wid, wv, fv = empty.transform(np.zeros((5, 32), np.uint8), 4)
    assert empty.size() == 0 and len(wid) == 0 and len(fv.node_ids) == 0
    empty.close()
    parent, leaf, desc, weight = arrays
    bad = parent.copy(); bad[1] = 3                                            # a child line before its parent's
    with pytest.raises(FrontendError):
        ORBVocabulary.from_arrays(ctx, 4, 3, bad, leaf, desc, weight)
    # all words stopped: empty vectors
    z = ORBVocabulary.from_arrays(ctx, 4, 3, parent, leaf, desc, np.zeros_like(weight))
    wid, wv, fv = z.transform(vocabulary_queries(desc, leaf, 50), 1)
    assert len(wid) == 0 and len(fv.node_ids) == 0
    z.close()


def test_descend_device_batch(orbvoc_shape, ctx):
    """Device-resident descriptors of a whole batch (what orbx_extract_batch_device leaves in HBM)."""
    import ctypes as C
    import torch
    from orb_slam_b200._lib import check, lib
    (parent, leaf, desc, weight), g, o = orbvoc_shape
    q = vocabulary_queries(desc, leaf, 20000, seed=8, max_flips=60)
    dq = torch.from_numpy(q).cuda()
    out = torch.zeros((len(q), 2), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    check(lib().orbm_vocab_descend_device(ctx.h, g._h, dq.data_ptr(), len(q), 4, out.data_ptr()), "descend_device")
    check(lib().orbm_synchronize(ctx.h), "sync")
    ow, onid, _ = o.descend(q, 4)
    res = out.cpu().numpy().view(np.uint32)
    assert np.array_equal(res[:, 0], ow) and np.array_equal(res[:, 1], onid)


@pytest.mark.parametrize("npoints,seed,max_obs", [(1, 1, 3), (200, 2, 12), (5000, 3, 9), (300, 4, 64)])
def test_distinctive_descriptors(ctx, npoints, seed, max_obs):
    from orb_slam_b200.vocabulary import compute_distinctive_descriptors
    D, start = observation_lists(npoints, seed=seed, max_obs=max_obs)
    best = compute_distinctive_descriptors(ctx, D, start)
    assert np.array_equal(best, oracle_distinctive_descriptors(D, start))
    sizes = np.diff(start)
    assert ((best >= 0) == (sizes > 0)).all() and (best < np.maximum(sizes, 1)).all()


def test_distinctive_descriptors_ties_and_empty(ctx):
    from orb_slam_b200.vocabulary import compute_distinctive_descriptors
    D = np.zeros((40, 32), np.uint8)                      # identical descriptors: every median is 0, the first row wins
    start = np.array([0, 0, 7, 40, 40], np.int32)         # empty, 7, 33 (> one warp), empty
    best = compute_distinctive_descriptors(ctx, D, start)
    assert best.tolist() == [-1, 0, 0, -1]
    assert np.array_equal(best, oracle_distinctive_descriptors(D, start))
    assert len(compute_distinctive_descriptors(ctx, np.zeros((0, 32), np.uint8), np.array([0], np.int32))) == 0
