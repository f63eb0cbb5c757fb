"""Synthetic grayscale frames for parity tests and benchmarks.

There are no datasets on either machine, so every configuration in
BASELINE.json runs on frames produced here: a counter-based generator
(splitmix64, seed = 0x0B5EED00 + frame_index) that sums three octaves of value
noise, paints 64 axis-aligned / rotated rectangles (strong corners at all
pyramid scales), squeezes one quadrant to +-10 gray levels (forces the
iniThFAST -> minThFAST fallback, reference src/ORBextractor.cc:982-987) and
flattens one stripe to gray 128 (forces empty cells and empty quadtree nodes).
Pure numpy; identical bytes on every machine.
"""
from __future__ import annotations

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)
BASE_SEED = 0x0B5EED00


def splitmix64(x: np.ndarray) -> np.ndarray:
    """Vectorised splitmix64 finaliser on uint64 arrays."""
    with np.errstate(over="ignore"):
        x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = x
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def _hash_grid(seed: int, salt: int, ny: int, nx: int) -> np.ndarray:
    """uniform [0,1) float32 lattice, counter = (seed, salt, iy, ix)"""
    iy, ix = np.mgrid[0:ny, 0:nx].astype(np.uint64)
    with np.errstate(over="ignore"):
        key = (np.uint64(seed) * np.uint64(0x100000001B3) + np.uint64(salt)) & _M64
        c = splitmix64(key ^ (iy << np.uint64(32)) ^ ix)
        c = splitmix64(c)
    return ((c >> np.uint64(40)).astype(np.float32)) * np.float32(1.0 / (1 << 24))


def _rand_stream(seed: int, salt: int, n: int) -> np.ndarray:
    return _hash_grid(seed, salt, 1, n)[0]


def _value_noise(seed: int, salt: int, h: int, w: int, cell: int) -> np.ndarray:
    ny, nx = h // cell + 2, w // cell + 2
    g = _hash_grid(seed, salt, ny, nx)
    ys = np.arange(h, dtype=np.float32) / np.float32(cell)
    xs = np.arange(w, dtype=np.float32) / np.float32(cell)
    y0 = ys.astype(np.int32); x0 = xs.astype(np.int32)
    fy = (ys - y0)[:, None]; fx = (xs - x0)[None, :]
    fy = fy * fy * (3 - 2 * fy); fx = fx * fx * (3 - 2 * fx)
    g00 = g[y0][:, x0]; g01 = g[y0][:, x0 + 1]; g10 = g[y0 + 1][:, x0]; g11 = g[y0 + 1][:, x0 + 1]
    top = g00 + (g01 - g00) * fx
    bot = g10 + (g11 - g10) * fx
    return top + (bot - top) * fy


def synth_frame(w: int, h: int, index: int = 0, *, shift: tuple[float, float] = (0.0, 0.0)) -> np.ndarray:
    """One u8 frame of size (h, w).  `shift` translates the whole scene by an
    integer number of pixels (frame-to-frame motion for the matcher configs)."""
    seed = BASE_SEED + int(index)
    sx, sy = int(round(shift[0])), int(round(shift[1]))
    pad = 64
    H, W = h + 2 * pad, w + 2 * pad
    im = np.full((H, W), 128.0, np.float32)
    for salt, (cell, amp) in enumerate(((4, 40.0), (16, 30.0), (64, 20.0))):
        im += (_value_noise(seed, salt, H, W, cell) - 0.5) * np.float32(2 * amp)
    r = _rand_stream(seed, 100, 64 * 6).reshape(64, 6)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    for k in range(64):
        cx, cy = r[k, 0] * W, r[k, 1] * H
        hw, hh = 6 + r[k, 2] * 70, 6 + r[k, 3] * 50
        gray = 20 + r[k, 4] * 215
        if k % 2:
            ang = r[k, 5] * np.pi
            ca, sa = np.float32(np.cos(ang)), np.float32(np.sin(ang))
            x0, x1 = int(max(cx - hw - hh, 0)), int(min(cx + hw + hh + 1, W))
            y0, y1 = int(max(cy - hw - hh, 0)), int(min(cy + hw + hh + 1, H))
            if x1 <= x0 or y1 <= y0:
                continue
            dx = xx[y0:y1, x0:x1] - cx; dy = yy[y0:y1, x0:x1] - cy
            m = (np.abs(dx * ca + dy * sa) <= hw) & (np.abs(-dx * sa + dy * ca) <= hh)
            im[y0:y1, x0:x1][m] = gray
        else:
            x0, x1 = int(max(cx - hw, 0)), int(min(cx + hw, W))
            y0, y1 = int(max(cy - hh, 0)), int(min(cy + hh, H))
            im[y0:y1, x0:x1] = gray
    # low-contrast quadrant (bottom-right) and flat stripe, anchored to the scene
    qy, qx = pad + h // 2, pad + w // 2
    im[qy:, qx:] = 128 + (im[qy:, qx:] - 128) * np.float32(10.0 / 128.0)
    sw = max(64, w // 12)
    x_s = pad + int(w * 0.30)
    im[:, x_s:x_s + sw] = 128
    im = np.clip(np.rint(im), 0, 255).astype(np.uint8)
    y0, x0 = pad - sy, pad - sx
    return np.ascontiguousarray(im[y0:y0 + h, x0:x0 + w])


def synth_batch(w: int, h: int, n: int, start: int = 0) -> np.ndarray:
    out = np.empty((n, h, w), np.uint8)
    for i in range(n):
        out[i] = synth_frame(w, h, start + i)
    return out


def synth_descriptors(n: int, seed: int) -> np.ndarray:
    """n x 32 uniform random bytes (256-bit descriptors), counter-based."""
    words = (n * 32 + 7) // 8
    idx = np.arange(words, dtype=np.uint64)
    with np.errstate(over="ignore"):
        v = splitmix64(splitmix64(np.uint64(seed)) ^ idx)
    return v.view(np.uint8)[: n * 32].reshape(n, 32).copy()


def synth_vocabulary(k: int = 10, L: int = 6, seed: int = 0xB0C, stop_fraction: float = 0.01):
    """A synthetic ORB vocabulary with the shape of Vocabulary/ORBvoc.txt (k = 10, L = 6: 1 111 111 nodes, 10^6
    words; the real file is absent from the reference tree).  Returns (parent, is_leaf, desc, weight) indexed by
    node id, node 0 = root, ids level by level so a parent precedes its children.  A child's descriptor is its
    parent's with a level-dependent number of random bit flips (cluster centres spread out near the root, tight
    near the leaves); leaf weights look like idf values, `stop_fraction` of them 0 (stopped words)."""
    rng = np.random.default_rng(seed)
    counts = [k ** l for l in range(L + 1)]
    n = sum(counts)
    parent = np.zeros(n, np.int32)
    desc = np.zeros((n, 32), np.uint8)
    first = np.cumsum([0] + counts)           # first id of each level
    for l in range(1, L + 1):
        ids = np.arange(first[l], first[l + 1])
        par = first[l - 1] + (ids - first[l]) // k
        parent[ids] = par
        if l == 1:
            desc[ids] = rng.integers(0, 256, (len(ids), 32), dtype=np.uint8)
        else:
            p_flip = 0.5 * (0.55 ** (l - 1))              # per-bit flip probability
            for c0 in range(0, len(ids), 1 << 16):                # chunks: 256 random draws per node
                c = slice(c0, c0 + (1 << 16))
                flips = np.packbits(rng.random((len(ids[c]), 256), dtype=np.float32) < p_flip, axis=1)
                desc[ids[c]] = desc[par[c]] ^ flips
    is_leaf = np.zeros(n, np.uint8)
    is_leaf[first[L]:] = 1
    weight = np.zeros(n, np.float64)
    nl = counts[L]
    w = np.log(rng.uniform(1.5, 4000.0, nl))
    w[rng.random(nl) < stop_fraction] = 0.0
    weight[first[L]:] = w
    return parent, is_leaf, desc, weight


def vocabulary_text(k, L, parent, is_leaf, desc, weight, scoring=0, weighting=0, trailing_newline=False) -> str:
    """The ORBvoc.txt text format (what TemplatedVocabulary::saveToTextFile writes / loadFromTextFile reads)."""
    lines = [f"{k} {L}  {scoring} {weighting}"]
    for i in range(1, len(parent)):
        lines.append(f"{int(parent[i])} {int(is_leaf[i])} " + " ".join(str(int(b)) for b in desc[i]) + f" {float(weight[i])!r}")
    return "\n".join(lines) + ("\n" if trailing_newline else "")
