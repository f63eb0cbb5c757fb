"""CPU checks of the SearchByBoW oracle (oracle/orb_oracle_bow.cc) against an independent pure-Python replay of
src/ORBmatcher.cc:159-288 / :590-723 on small scenes (the reference ships no fixtures for the matcher)."""
import numpy as np
import pytest

from tests import scenes
from tests.oracle_lib import OracleMatcher

POP = np.array([bin(i).count("1") for i in range(256)], np.int32)


def three_maxima(counts):
    max1 = max2 = max3 = 0
    i1 = i2 = i3 = -1
    for i, s in enumerate(counts):
        if s > max1:
            max3, max2, max1 = max2, max1, s; i3, i2, i1 = i2, i1, i
        elif s > max2:
            max3, max2 = max2, s; i3, i2 = i2, i
        elif s > max3:
            max3 = s; i3 = i
    if max2 < np.float32(0.1) * np.float32(max1):
        i2 = i3 = -1
    elif max3 < np.float32(0.1) * np.float32(max1):
        i3 = -1
    return i1, i2, i3


def py_bow(variant, s, nnratio, ori):
    fv1 = {int(k): s["fv1"].items[s["fv1"].start[i]:s["fv1"].start[i + 1]] for i, k in enumerate(s["fv1"].node_ids)}
    fv2 = {int(k): s["fv2"].items[s["fv2"].start[i]:s["fv2"].start[i + 1]] for i, k in enumerate(s["fv2"].node_ids)}
    n_out = len(s["desc2"]) if variant == 0 else len(s["desc1"])
    match = np.full(n_out, -1, np.int32)
    taken = np.zeros(len(s["desc2"]), bool)
    hist = [[] for _ in range(30)]
    nm = 0
    for node in sorted(set(fv1) & set(fv2)):
        for i1 in fv1[node]:
            if not s["valid1"][i1]:
                continue
            b1, b2, bi = 256, 256, -1
            for i2 in fv2[node]:
                if variant == 0 and match[i2] >= 0:
                    continue
                if variant == 1 and (taken[i2] or not s["valid2"][i2]):
                    continue
                d = int(POP[s["desc1"][i1] ^ s["desc2"][i2]].sum())
                if d < b1:
                    b2, b1, bi = b1, d, int(i2)
                elif d < b2:
                    b2 = d
            ok = b1 <= 50 if variant == 0 else b1 < 50
            if ok and np.float32(b1) < np.float32(nnratio) * np.float32(b2):
                if variant == 0:
                    match[bi] = i1
                else:
                    match[i1] = bi; taken[bi] = True
                if ori:
                    rot = np.float32(s["angle1"][i1]) - np.float32(s["angle2"][bi])
                    if rot < 0:
                        rot = np.float32(rot + np.float32(360.0))
                    v = float(np.float32(rot * np.float32(1.0 / 30)))
                    b = int(np.floor(v + 0.5))                                   # C round(): ties away from zero (v >= 0)
                    if b == 30:
                        b = 0
                    hist[b].append(bi if variant == 0 else int(i1))
                nm += 1
    if ori:
        keep = three_maxima([len(h) for h in hist])
        for i, h in enumerate(hist):
            if i in keep:
                continue
            for idx in h:
                match[idx] = -1; nm -= 1
    return nm, match


@pytest.mark.parametrize("seed", range(6))
@pytest.mark.parametrize("variant", [0, 1])
def test_oracle_matches_python_replay(seed, variant):
    s = scenes.bow_scene(seed=100 + seed, n1=150 + 20 * seed, n2=170, n_nodes=5 + seed, ties=10 * seed, flips=20)
    o = OracleMatcher(0.75, seed % 2 == 0)
    if variant == 0:
        n, m = o.SearchByBoWFrame(s["desc1"], s["angle1"], s["valid1"], s["fv1"], s["desc2"], s["angle2"], s["fv2"])
    else:
        n, m = o.SearchByBoWKeyFrames(s["desc1"], s["angle1"], s["valid1"], s["fv1"], s["desc2"], s["angle2"], s["valid2"], s["fv2"])
    pn, pm = py_bow(variant, s, 0.75, seed % 2 == 0)
    assert n == pn
    np.testing.assert_array_equal(m, pm)
    assert n == int((m >= 0).sum())


def test_oracle_empty_inputs():
    s = scenes.bow_scene(seed=1, n1=0, n2=10)
    n, m = OracleMatcher().SearchByBoWFrame(s["desc1"], s["angle1"], s["valid1"], s["fv1"], s["desc2"], s["angle2"], s["fv2"])
    assert n == 0 and np.all(m == -1) and len(m) == 10
