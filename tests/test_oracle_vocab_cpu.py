"""CPU tests of the vocabulary oracle (oracle/orb_oracle_vocab.cc): against the golden vectors produced by the
reference's own DBoW2 (tests/golden/vocab_ref.npz, generator make_vocab_golden.py), against that build directly when
it is present (oracle/_ref/libdbow2_ref.so), and the distinctive-descriptor selection against a plain replay."""
import os

import numpy as np
import pytest

from orb_slam_b200.synth import synth_vocabulary, vocabulary_text
from tests.oracle_lib import OracleVocabulary, RefVocabulary, have_dbow2_ref, oracle_distinctive_descriptors
from tests.scenes import observation_lists, unbalanced_vocabulary, vocabulary_queries

GOLD = os.path.join(os.path.dirname(__file__), "golden", "vocab_ref.npz")


def golden_cases():
    z = np.load(GOLD)
    names = sorted({k.split("/")[0] for k in z.files})
    return z, names


def fv_from_flat(ids, start, items):
    return {int(ids[j]): items[start[j]:start[j + 1]].tolist() for j in range(len(ids))}


@pytest.mark.parametrize("name", golden_cases()[1])
def test_oracle_matches_reference_dbow2_golden(name):
    z, _ = golden_cases()
    k, L, scoring, weighting, levelsup = (int(v) for v in z[name + "/meta"])
    voc = OracleVocabulary.from_arrays(k, L, z[name + "/parent"], z[name + "/leaf"], z[name + "/desc"], z[name + "/weight"],
                                       scoring, weighting)
    wid, wv, fv = voc.transform(z[name + "/queries"], levelsup)
    assert np.array_equal(wid, z[name + "/word_ids"])
    assert np.array_equal(wv, z[name + "/word_vals"])                 # doubles, bit for bit
    assert fv == fv_from_flat(z[name + "/node_ids"], z[name + "/node_start"], z[name + "/items"])


@pytest.mark.skipif(not have_dbow2_ref(), reason="oracle/_ref/libdbow2_ref.so not built (needs /root/reference)")
@pytest.mark.parametrize("weighting,scoring", [(0, 0), (1, 0), (2, 1), (3, 5), (1, 5), (0, 2)])
def test_oracle_matches_reference_dbow2_live(tmp_path, weighting, scoring):
    parent, leaf, desc, weight = synth_vocabulary(9, 4, seed=20 + weighting, stop_fraction=0.03)
    path = tmp_path / "voc.txt"
    path.write_text(vocabulary_text(9, 4, parent, leaf, desc, weight, scoring, weighting))
    ref = RefVocabulary(path)
    voc = OracleVocabulary.loadFromTextFile(path)                     # the oracle's own text loader
    assert voc.size() == ref.size() == 9 ** 4
    q = vocabulary_queries(desc, leaf, 1500, seed=weighting * 7 + scoring)
    for levelsup in (0, 2, 4):
        a, b = voc.transform(q, levelsup), ref.transform(q, levelsup)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2] == b[2]


def test_text_loader_skips_blank_lines(tmp_path):
    parent, leaf, desc, weight = unbalanced_vocabulary(5, 3, seed=2)
    txt = vocabulary_text(5, 3, parent, leaf, desc, weight)
    (tmp_path / "a.txt").write_text(txt)
    (tmp_path / "b.txt").write_text(txt + "\n\n")
    a, b = OracleVocabulary.loadFromTextFile(tmp_path / "a.txt"), OracleVocabulary.loadFromTextFile(tmp_path / "b.txt")
    assert a.nodes() == b.nodes() == len(parent) and a.size() == int(leaf.sum())
    q = vocabulary_queries(desc, leaf, 200)
    ra, rb = a.transform(q, 1), b.transform(q, 1)
    assert np.array_equal(ra[0], rb[0]) and np.array_equal(ra[1], rb[1]) and ra[2] == rb[2]
    c = OracleVocabulary.from_arrays(5, 3, parent, leaf, desc, weight)
    rc = c.transform(q, 1)
    assert np.array_equal(ra[0], rc[0]) and np.array_equal(ra[1], rc[1]) and ra[2] == rc[2]


def test_descend_first_minimum_wins():
    # two identical sibling centres: the scan keeps the first one (strict '<', TemplatedVocabulary.h:1244)
    parent = np.array([0, 0, 0, 0], np.int32); leaf = np.array([0, 1, 1, 1], np.uint8)
    desc = np.zeros((4, 32), np.uint8); desc[1] = 0xFF; desc[2] = 0x0F; desc[3] = 0x0F
    weight = np.array([0, 1.0, 2.0, 3.0])
    voc = OracleVocabulary.from_arrays(3, 1, parent, leaf, desc, weight)
    w, nid, wt = voc.descend(np.full((1, 32), 0x0F, np.uint8), 0)
    assert (w[0], nid[0], wt[0]) == (1, 2, 2.0)


def test_distinctive_descriptors_against_replay():
    D, start = observation_lists(60, seed=4)
    best = oracle_distinctive_descriptors(D, start)
    bits = np.unpackbits(D, axis=1).astype(np.int32)
    for p in range(len(start) - 1):
        d = bits[start[p]:start[p + 1]]
        n = len(d)
        if n == 0:
            assert best[p] == -1
            continue
        dist = (d[:, None, :] != d[None, :, :]).sum(axis=2)
        med = np.sort(dist, axis=1)[:, int(0.5 * (n - 1))]
        assert best[p] == int(np.argmin(med))                        # argmin = first minimum
