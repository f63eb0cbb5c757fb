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


def fuse_points(src: Frame, seed=0):
    """Map points seen from a neighbouring key frame `src` (pose identity) that LocalMapping::SearchInNeighbors would try
    to fuse into another key frame: keyframe_points plus the mean viewing direction (MapPoint::GetNormal) and the
    IsInKeyFrame flag."""
    rng = np.random.default_rng(seed + 1000)
    pts, mm, _ = keyframe_points(src, seed=seed, p_valid=0.9)
    w = pts["world"].astype(np.float32)
    normals = w / np.linalg.norm(w, axis=1, keepdims=True)
    normals = normals + rng.normal(0, 0.15, normals.shape)
    flip = rng.random(len(w)) < 0.05                     # a few seen from the other side: viewing-angle gate
    normals[flip] *= -1
    normals = (normals / np.linalg.norm(normals, axis=1, keepdims=True)).astype(np.float32)
    in_kf = (rng.random(len(w)) < 0.2).astype(np.uint8)
    return pts, mm, normals, in_kf


def triangulation_scene(seed=0, n=1500, extra=400, t=(0.4, 0.02, 0.05), stereo_frac=0.5, n_nodes=60, flips=6, outlier_frac=0.1):
    """Two key frames for ORBmatcher::SearchForTriangulation: key frame 1 at the origin, key frame 2 = translation t.
    Key frame 2 observes noisy reprojections of key frame 1's points (descriptors with up to `flips` flipped bits; a
    share displaced off the epipolar line) plus unrelated keypoints.  F12 as LocalMapping::ComputeF12
    (src/LocalMapping.cc:651-668) with R12 = I, t12 = -t."""
    from orb_slam_b200._lib import KEYPOINT_DTYPE
    from orb_slam_b200.matcher import FeatureVector, KeyFrameView
    rng = np.random.default_rng(seed)
    K = KITTI_K.astype(np.float32)
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    sf = scale_factors(8, 1.2); sig2 = (sf * sf).astype(np.float32)
    k1 = np.zeros(n, KEYPOINT_DTYPE)
    k1["x"] = rng.uniform(20, 1220, n); k1["y"] = rng.uniform(20, 355, n)
    k1["octave"] = rng.integers(0, 8, n); k1["angle"] = rng.uniform(0, 360, n); k1["size"] = 31
    z = rng.uniform(4, 40, n).astype(np.float32)
    X = np.stack([(k1["x"] - cx) * z / fx, (k1["y"] - cy) * z / fy, z], axis=1) + np.asarray(t, np.float32)
    n2 = n + extra
    k2 = np.zeros(n2, KEYPOINT_DTYPE)
    perm = rng.permutation(n2)[:n]
    k2["x"] = rng.uniform(20, 1220, n2); k2["y"] = rng.uniform(20, 355, n2)
    k2["octave"] = rng.integers(0, 8, n2); k2["angle"] = rng.uniform(0, 360, n2); k2["size"] = 31
    k2["x"][perm] = fx * X[:, 0] / X[:, 2] + cx + rng.normal(0, 0.4, n)
    k2["y"][perm] = fy * X[:, 1] / X[:, 2] + cy + rng.normal(0, 0.4, n)
    off = rng.random(n) < outlier_frac
    k2["y"][perm[off]] += rng.choice([-1, 1], off.sum()) * rng.uniform(8, 30, off.sum())
    k2["octave"][perm] = np.clip(k1["octave"] + rng.integers(-1, 2, n), 0, 7)
    k2["angle"][perm] = (k1["angle"] - 12 + rng.normal(0, 3, n)) % 360
    d1 = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    d2 = rng.integers(0, 256, (n2, 32), dtype=np.uint8)
    for a, b in zip(range(n), perm):
        d = d1[a].copy()
        for bit in rng.integers(0, 256, rng.integers(0, flips + 1)):
            d[bit >> 3] ^= np.uint8(1 << (bit & 7))
        d2[b] = d
    if n >= 80:
        d2[perm[:40]] = d2[perm[40:80]]                   # equal distances inside a node: ties go to the LAST candidate

    def node_of(d):
        return ((d[:, 1].astype(np.uint32) >> 2) % np.uint32(n_nodes)) * 2 + 5

    def featvec(d):
        nodes = {}
        for i, nd in enumerate(node_of(d)):
            nodes.setdefault(int(nd), []).append(i)
        return FeatureVector(nodes)

    def uright(k):
        u = np.full(len(k), -1, np.float32)
        st = rng.random(len(k)) < stereo_frac
        u[st] = k["x"][st] - rng.uniform(2, 60, st.sum()).astype(np.float32)
        return u

    kf1 = KeyFrameView(k1, d1, uright(k1), (rng.random(n) < 0.3).astype(np.uint8), featvec(d1))
    kf2 = KeyFrameView(k2, d2, uright(k2), (rng.random(n2) < 0.3).astype(np.uint8), featvec(d2))
    tt = np.asarray(t, np.float64); t12 = -tt
    t12x = np.array([[0, -t12[2], t12[1]], [t12[2], 0, -t12[0]], [-t12[1], t12[0], 0]])
    Kd = K.astype(np.float64)
    F12 = (np.linalg.inv(Kd).T @ t12x @ np.linalg.inv(Kd)).astype(np.float32)
    return dict(kf1=kf1, kf2=kf2, Cw1=np.zeros(3, np.float32), R2w=np.eye(3, dtype=np.float32), t2w=np.asarray(t, np.float32),
                K2=K, sf=sf, sig2=sig2, F12=F12, true12=perm, off=off)


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


def sim3_terms(s12, R12, t12):
    """sR12, sR21, t21 of src/ORBmatcher.cc:1187-1189 with OpenCV's arithmetic written out: a scaled Mat is a float product by
    float(alpha), the 3x3 * 3x1 product sums float products and the alpha = -1 of the unary minus is applied in double."""
    s12 = np.float32(s12); R12 = np.asarray(R12, np.float32); t12 = np.asarray(t12, np.float32)
    sR12 = (R12 * s12).astype(np.float32)
    sR21 = (R12.T * np.float32(1.0 / np.float64(s12))).astype(np.float32)
    t21 = np.zeros(3, np.float32)
    for r in range(3):
        acc = np.float32(sR21[r, 0] * t12[0])
        acc = np.float32(acc + np.float32(sR21[r, 1] * t12[1]))
        acc = np.float32(acc + np.float32(sR21[r, 2] * t12[2]))
        t21[r] = np.float32(np.float64(acc) * -1.0)
    return sR12, sR21, t21


def sim3_scene(seed=0, n=1500, extra=300, s12=1.08, flips=5):
    """Two key frames of a loop candidate for ORBmatcher::SearchBySim3.  Key frame 1 at the origin of its map, key frame 2 at
    pose T2w in a drifted map; camera-1 coordinates = s12 R12 camera-2 coordinates + t12.  n points are seen by both (keypoints
    at each other's projections plus noise, descriptors up to `flips` bits apart, octaves compatible with PredictScale);
    `extra` unrelated keypoints with map points on each side; some slots without point, bad points, already-matched ones."""
    from orb_slam_b200._lib import KEYPOINT_DTYPE
    rng = np.random.default_rng(seed)
    K = KITTI_K.astype(np.float64)
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    sf = scale_factors(8, 1.2).astype(np.float64)
    a = 0.01
    R12 = np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]], np.float32)
    t12 = np.array([0.3, -0.05, 0.4], np.float32)
    sR12, sR21, t21 = sim3_terms(s12, R12, t12)
    b = -0.02
    R2w = np.array([[np.cos(b), -np.sin(b), 0], [np.sin(b), np.cos(b), 0], [0, 0, 1]], np.float32)
    T2w = np.eye(4, dtype=np.float32); T2w[:3, :3] = R2w; T2w[:3, 3] = [5.0, -1.0, 2.0]
    N = n + extra

    def keys(N):
        k = np.zeros(N, KEYPOINT_DTYPE)
        k["x"] = rng.uniform(20, 1220, N); k["y"] = rng.uniform(20, 355, N)
        k["octave"] = rng.integers(0, 8, N); k["angle"] = rng.uniform(0, 360, N); k["size"] = 31
        return k
    k1, k2 = keys(N), keys(N)
    d1 = rng.integers(0, 256, (N, 32), dtype=np.uint8); d2 = rng.integers(0, 256, (N, 32), dtype=np.uint8)
    z = rng.uniform(5, 40, N)
    P1 = np.stack([(k1["x"] - cx) * z / fx, (k1["y"] - cy) * z / fy, z], axis=1)          # camera-1 = world-1 coordinates
    i2 = rng.permutation(N)[:n]                                                          # shared point i <-> keypoint i2[i] of key frame 2
    Pc2 = P1[:n] @ sR21.astype(np.float64).T + t21                                        # in camera 2
    ok = Pc2[:, 2] > 1.0
    k2["x"][i2] = np.where(ok, fx * Pc2[:, 0] / Pc2[:, 2] + cx + rng.normal(0, 0.5, n), k2["x"][i2])
    k2["y"][i2] = np.where(ok, fy * Pc2[:, 1] / Pc2[:, 2] + cy + rng.normal(0, 0.5, n), k2["y"][i2])
    d2[i2] = d1[:n]
    for _ in range(flips):
        d2[i2, rng.integers(0, 32, n)] ^= (1 << rng.integers(0, 8, n)).astype(np.uint8)
    k2["octave"][i2] = np.clip(k1["octave"][:n] + rng.integers(-1, 2, n), 0, 7)
    # key frame 2's own map points: camera-2 coordinates back through T2w
    z2 = rng.uniform(5, 40, N)
    Pc2_all = np.stack([(k2["x"] - cx) * z2 / fx, (k2["y"] - cy) * z2 / fy, z2], axis=1)
    Pc2_all[i2] = np.where(ok[:, None], Pc2 + rng.normal(0, 0.01, Pc2.shape), Pc2_all[i2])
    W2 = (Pc2_all - T2w[:3, 3].astype(np.float64)) @ R2w.astype(np.float64)               # R2w^T (p - t2w)
    Pc1_of2 = Pc2_all @ sR12.astype(np.float64).T + t12                                   # where key frame 2's points fall in camera 1

    def points(world, desc, koct_other, dist_other):
        p = np.zeros(N, MAPPOINT_DTYPE)
        p["world"] = world.astype(np.float32); p["descriptor"] = desc
        p["valid"] = rng.random(N) < 0.9; p["bad"] = rng.random(N) < 0.04
        # mfMaxDistance such that PredictScale in the other key frame gives the other keypoint's octave or one above
        mx = dist_other * sf[1] ** (koct_other + rng.uniform(-0.9, 0.9, N))
        mm = np.stack([mx / sf[-1], mx], axis=1).astype(np.float32)
        return p, mm
    oct2_for1 = k2["octave"].copy().astype(np.float64); oct2_for1[:] = rng.integers(0, 8, N)
    o2 = np.full(N, 3.0); o2[:n] = k2["octave"][i2]
    dist_1in2 = np.linalg.norm(np.concatenate([Pc2, P1[n:] @ sR21.astype(np.float64).T + t21]), axis=1)
    pts1, mm1 = points(P1, d1, o2, dist_1in2)
    o1 = np.full(N, 3.0); o1[i2] = k1["octave"][:n]
    pts2, mm2 = points(W2, d2, o1, np.linalg.norm(Pc1_of2, axis=1))
    kf1 = make_frame(k1, d1, 1241, 376)
    kf2 = make_frame(k2, d2, 1241, 376, Tcw=T2w)
    skip1 = (pts1["bad"].astype(bool) | (rng.random(N) < 0.1)).astype(np.uint8)
    skip2 = (pts2["bad"].astype(bool) | (rng.random(N) < 0.1)).astype(np.uint8)
    return dict(kf1=kf1, kf2=kf2, pts1=pts1, pts2=pts2, mm1=mm1, mm2=mm2, skip1=skip1, skip2=skip2, s12=np.float32(s12), R12=R12,
                t12=t12, sR12=sR12, sR21=sR21, t21=t21, i2=i2)


# ---------------------------------------------------------------- vocabulary scenes (ComputeBoW)
def unbalanced_vocabulary(k=7, L=4, seed=3, p_leaf=0.2, stop_fraction=0.1, min_leaf_level=1):
    """Small vocabulary in the id order HKmeansStep creates (a node's children get consecutive ids, then each child's
    subtree follows: Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:640-760), with ragged child counts and leaves at
    different depths (none above `min_leaf_level`: transform() leaves the FeatureVector node id uninitialised when a leaf is
    shallower than level L - levelsup).  Returns (parent, is_leaf, desc, weight)."""
    rng = np.random.default_rng(seed)
    parent, leaf, desc, weight = [0], [0], [np.zeros(32, np.uint8)], [0.0]

    def grow(node, level):
        nchild = int(rng.integers(2, k + 1))
        ids = []
        for _ in range(nchild):
            ids.append(len(parent))
            parent.append(node); leaf.append(0); weight.append(0.0)
            if level == 1:
                desc.append(rng.integers(0, 256, 32, dtype=np.uint8))
            else:
                flips = np.packbits(rng.random(256) < 0.5 * 0.55 ** (level - 1))
                desc.append(desc[node] ^ flips)
        for c in ids:
            if level < L and (level < min_leaf_level or rng.random() > p_leaf):
                grow(c, level + 1)
            else:
                leaf[c] = 1
                weight[c] = 0.0 if rng.random() < stop_fraction else float(np.log(rng.uniform(1.5, 4000.0)))

    grow(0, 1)
    return (np.array(parent, np.int32), np.array(leaf, np.uint8), np.stack(desc).astype(np.uint8),
            np.array(weight, np.float64))


def vocabulary_queries(desc, is_leaf, n, seed=5, max_flips=40):
    """n query descriptors: vocabulary leaves with 0..max_flips random bit flips (exact ties and near-ties between
    sibling centres occur), every fourth one uniformly random."""
    rng = np.random.default_rng(seed)
    leaves = np.flatnonzero(is_leaf)
    q = desc[rng.choice(leaves, n)].copy()
    for i in range(n):
        if i % 4 == 3:
            q[i] = rng.integers(0, 256, 32, dtype=np.uint8)
        else:
            bits = rng.choice(256, int(rng.integers(0, max_flips + 1)), replace=False)
            for b in bits:
                q[i, b >> 3] ^= np.uint8(1 << (b & 7))
    return q


def observation_lists(npoints, seed=9, max_obs=12, big=(40, 77, 150)):
    """Descriptor lists of map points for ComputeDistinctiveDescriptors: a true descriptor per point observed with
    noise 1..max_obs times (plus a few long lists, an empty one and duplicates that force median ties)."""
    rng = np.random.default_rng(seed)
    descs, start = [], [0]
    sizes = [int(rng.integers(1, max_obs + 1)) for _ in range(npoints)]
    for j, b in enumerate(big):
        if j < npoints:
            sizes[(j * 7 + 3) % npoints] = b
    if npoints > 5:
        sizes[5] = 0
    for p, n in enumerate(sizes):
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        for o in range(n):
            d = base.copy()
            if not (p % 3 == 0 and o % 2 == 1):          # every other observation of every third point is an exact duplicate
                for b in rng.choice(256, int(rng.integers(0, 30)), replace=False):
                    d[b >> 3] ^= np.uint8(1 << (b & 7))
            descs.append(d)
        start.append(start[-1] + n)
    D = np.stack(descs).astype(np.uint8) if descs else np.zeros((0, 32), np.uint8)
    return D, np.array(start, np.int32)
