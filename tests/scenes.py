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
