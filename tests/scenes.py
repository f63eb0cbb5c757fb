"""Synthetic matcher scenes shared by the CPU and GPU tests: frames built from oracle-extracted
keypoints, map points back-projected at random depths, KITTI-like intrinsics."""
from __future__ import annotations

import numpy as np

from orb_slam_b200.matcher import MAPPOINT_DTYPE, Frame
from orb_slam_b200.synth import synth_frame

KITTI_K = np.array([[718.856, 0, 607.1928], [0, 718.856, 185.2157], [0, 0, 1]], np.float32)   # KITTI00-02.yaml:8-11
KITTI_BF = 386.1448


def extract_pair(extractor_factory, w, h, nf, idx, shift):
    ex = extractor_factory(nf, 1.2, 8, 20, 7)
    a = synth_frame(w, h, idx)
    b = synth_frame(w, h, idx, shift=shift)
    ka, da = ex(a)
    kb, db = ex(b)
    return (ka, da), (kb, db)


def scale_factors(n=8, sf=1.2):
    out = [np.float32(1.0)]
    for _ in range(1, n):
        out.append(np.float32(np.float64(out[-1]) * np.float64(np.float32(sf))))
    return np.array(out, np.float32)


def make_frame(k, d, w, h, *, stereo=False, Tcw=None, seed=0):
    rng = np.random.default_rng(seed)
    ur = None
    if stereo:
        ur = (k["x"] - rng.uniform(3, 60, len(k))).astype(np.float32)
        ur[rng.random(len(k)) < 0.3] = -1
    return Frame(k, d, w, h, scale_factors(), u_right=ur, K=KITTI_K, bf=KITTI_BF, Tcw=Tcw)


def last_frame_points(last: Frame, seed=0, p_valid=0.8):
    """MapPoints of the last frame: keypoint i back-projected at a random depth (last pose = identity)."""
    rng = np.random.default_rng(seed)
    n = last.N
    pts = np.zeros(n, MAPPOINT_DTYPE)
    z = rng.uniform(5, 50, n).astype(np.float32)
    pts["world"][:, 0] = (last.mvKeysUn["x"] - last.cx) * z / last.fx
    pts["world"][:, 1] = (last.mvKeysUn["y"] - last.cy) * z / last.fy
    pts["world"][:, 2] = z
    pts["descriptor"] = last.mDescriptors
    flip = rng.integers(0, 32, n)                       # a few flipped bits so distances are not all zero
    pts["descriptor"][np.arange(n), flip] ^= rng.integers(0, 256, n).astype(np.uint8)
    pts["valid"] = rng.random(n) < p_valid
    pts["outlier"] = rng.random(n) < 0.05
    pts["observations"] = rng.integers(0, 4, n)
    return pts


def local_map_points(F: Frame, seed=0):
    rng = np.random.default_rng(seed)
    n = F.N
    pts = np.zeros(n, MAPPOINT_DTYPE)
    pts["descriptor"] = F.mDescriptors
    flip = rng.integers(0, 32, n)
    pts["descriptor"][np.arange(n), flip] ^= rng.integers(0, 256, n).astype(np.uint8)
    pts["valid"] = 1
    pts["bad"] = rng.random(n) < 0.05
    pts["trackInView"] = rng.random(n) < 0.8
    pts["trackProjX"] = F.mvKeysUn["x"] + rng.normal(0, 1.5, n).astype(np.float32)
    pts["trackProjY"] = F.mvKeysUn["y"] + rng.normal(0, 1.5, n).astype(np.float32)
    pts["trackProjXR"] = pts["trackProjX"] - rng.uniform(3, 60, n).astype(np.float32)
    pts["trackViewCos"] = rng.uniform(0.99, 1.0, n).astype(np.float32)
    pts["trackScaleLevel"] = np.clip(F.mvKeysUn["octave"] + rng.integers(0, 2, n), 0, 7)
    pts["observations"] = rng.integers(0, 4, n)
    return pts


def keyframe_points(kf: Frame, seed=0, p_valid=0.85):
    """Map points of a key frame for the relocalisation matcher: keypoint i back-projected at a random depth (key-frame
    pose = identity), with the scale-invariance distances MapPoint::UpdateNormalAndDepth would give it
    (src/MapPoint.cc:352-370: mfMaxDistance = dist * scaleFactor^level, mfMinDistance = mfMaxDistance / scaleFactor^(nLevels-1))."""
    rng = np.random.default_rng(seed)
    n = kf.N
    pts = np.zeros(n, MAPPOINT_DTYPE)
    z = rng.uniform(4, 60, n).astype(np.float32)
    pts["world"][:, 0] = (kf.mvKeysUn["x"] - kf.cx) * z / kf.fx
    pts["world"][:, 1] = (kf.mvKeysUn["y"] - kf.cy) * z / kf.fy
    pts["world"][:, 2] = z
    pts["descriptor"] = kf.mDescriptors
    flip = rng.integers(0, 32, n)
    pts["descriptor"][np.arange(n), flip] ^= rng.integers(0, 256, n).astype(np.uint8)
    pts["valid"] = rng.random(n) < p_valid
    pts["bad"] = rng.random(n) < 0.05
    dist = np.linalg.norm(pts["world"], axis=1).astype(np.float32)
    sf = kf.mvScaleFactors
    mx = (dist * sf[np.clip(kf.mvKeysUn["octave"], 0, len(sf) - 1)]).astype(np.float32)
    mn = (mx / sf[-1]).astype(np.float32)
    mm = np.stack([mn, mx], axis=1).astype(np.float32)
    already = (rng.random(n) < 0.2).astype(np.uint8)
    return pts, mm, already


def bow_scene(seed=0, n1=1500, n2=1600, *, n_nodes=64, p_common=0.7, flips=12, p_valid=0.8, rotation=25.0, ties=0):
    """Two descriptor sets with a synthetic vocabulary for ORBmatcher::SearchByBoW: side 2 holds noisy copies (up to
    `flips` flipped bits) of a share of side 1 plus unrelated descriptors; a feature's node is a hash of a few
    descriptor bits (so some true pairs land in different nodes, as with a real vocabulary tree); feature vectors
    list the features of a node in ascending index (DBoW2 appends them in feature order).  `ties` duplicates that
    many side-2 descriptors inside their node so equal distances occur.  Returns a dict of arrays + FeatureVectors."""
    from orb_slam_b200.matcher import FeatureVector
    rng = np.random.default_rng(seed)
    d1 = rng.integers(0, 256, (n1, 32), dtype=np.uint8)
    d2 = rng.integers(0, 256, (n2, 32), dtype=np.uint8)
    ncommon = int(min(n1, n2) * p_common)
    src = rng.permutation(n1)[:ncommon]
    dst = rng.permutation(n2)[:ncommon]
    for a, b in zip(src, dst):
        d = d1[a].copy()
        for bit in rng.integers(0, 256, rng.integers(0, flips + 1)):
            d[bit >> 3] ^= np.uint8(1 << (bit & 7))
        d2[b] = d
    for _ in range(ties):
        if n2 >= 2:
            a, b = rng.integers(0, n2, 2)
            d2[b] = d2[a]
    a1 = rng.uniform(0, 360, n1).astype(np.float32)
    a2 = rng.uniform(0, 360, n2).astype(np.float32)
    a2[dst] = ((a1[src] - np.float32(rotation) + rng.normal(0, 3, ncommon)) % 360).astype(np.float32)

    def node_of(d):
        return ((d[:, 0].astype(np.uint32) ^ (d[:, 7].astype(np.uint32) << 1)) % np.uint32(max(n_nodes, 1))) * 3 + 11

    def featvec(d, drop):
        nodes = {}
        for i, nd in enumerate(node_of(d)):
            if nd in drop:
                continue
            nodes.setdefault(int(nd), []).append(i)
        return FeatureVector(nodes)

    ids = np.unique(np.concatenate([node_of(d1), node_of(d2)])) if n1 + n2 else np.zeros(0, np.uint32)
    drop1 = set(int(x) for x in ids[rng.random(len(ids)) < 0.1])       # nodes missing on one side exercise lower_bound
    drop2 = set(int(x) for x in ids[rng.random(len(ids)) < 0.1])
    v1 = (rng.random(n1) < p_valid).astype(np.uint8)
    v2 = (rng.random(n2) < p_valid).astype(np.uint8)
    return dict(desc1=d1, desc2=d2, angle1=a1, angle2=a2, valid1=v1, valid2=v2, fv1=featvec(d1, drop1), fv2=featvec(d2, drop2))
