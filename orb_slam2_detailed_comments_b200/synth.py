"""Deterministic synthetic frames and descriptor sets.

The reference ships no data (its examples read TUM/KITTI/EuRoC from disk, Examples/Monocular/mono_tum.cc:70),
so every test and benchmark uses these seeded generators at the dataset shapes.
"""
from __future__ import annotations

import numpy as np

CONFIGS = {
    # name: (width, height, nFeatures)  -- 8 levels, scale 1.2, FAST 20/7 everywhere
    "tum": (640, 480, 1000),       # Examples/Monocular/TUM1.yaml:30-43
    "kitti": (1241, 376, 2000),    # Examples/Stereo/KITTI00-02.yaml:38-51
    "euroc": (752, 480, 1200),     # Examples/Stereo/EuRoC.yaml:93-106
    "4k": (3840, 2160, 8000),
}


def synth_frame(width: int, height: int, seed: int) -> np.ndarray:
    """One u8 frame: smooth background + W*H/1500 filled rotated rectangles, soft-blurred, + noise."""
    rng = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:height, 0:width].astype(np.float32)
    img = 96.0 + 48.0 * np.sin(xx / 97.0) + 32.0 * np.cos(yy / 61.0)
    nrect = (width * height) // 1500
    cx = rng.uniform(0, width, nrect)
    cy = rng.uniform(0, height, nrect)
    sw = rng.uniform(6, 80, nrect)
    sh = rng.uniform(6, 80, nrect)
    ang = np.deg2rad(rng.uniform(0, 180, nrect))
    grey = rng.randint(0, 256, nrect)
    for i in range(nrect):
        r = 0.5 * np.hypot(sw[i], sh[i]) + 1
        x0, x1 = max(int(cx[i] - r), 0), min(int(cx[i] + r) + 1, width)
        y0, y1 = max(int(cy[i] - r), 0), min(int(cy[i] + r) + 1, height)
        if x0 >= x1 or y0 >= y1:
            continue
        dx = xx[y0:y1, x0:x1] - cx[i]
        dy = yy[y0:y1, x0:x1] - cy[i]
        c, s = np.cos(ang[i]), np.sin(ang[i])
        u = dx * c + dy * s
        v = -dx * s + dy * c
        m = (np.abs(u) <= 0.5 * sw[i]) & (np.abs(v) <= 0.5 * sh[i])
        img[y0:y1, x0:x1][m] = grey[i]
    # separable Gaussian sigma 0.8 (radius 3), reflect border
    t = np.arange(-3, 4, dtype=np.float32)
    k = np.exp(-t * t / (2 * 0.8 * 0.8))
    k /= k.sum()
    p = np.pad(img, 3, mode="reflect")
    tmp = sum(k[i] * p[:, i:i + width] for i in range(7))
    img = sum(k[i] * tmp[i:i + height, :] for i in range(7))
    img = img + rng.uniform(-3, 3, img.shape).astype(np.float32)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def synth_batch(width: int, height: int, n: int, seed0: int = 1234, distinct: int | None = None) -> np.ndarray:
    """n frames (n, H, W) u8; seeds seed0+i. `distinct` < n generates that many and tiles them with a
    per-copy circular shift so every frame still differs (used by large throughput batches)."""
    d = n if distinct is None else min(distinct, n)
    base = [synth_frame(width, height, seed0 + i) for i in range(d)]
    out = np.empty((n, height, width), np.uint8)
    for i in range(n):
        f = base[i % d]
        rep = i // d
        out[i] = f if rep == 0 else np.roll(f, (7 * rep, 13 * rep), axis=(0, 1))
    return out


def synth_stereo_pair(width: int, height: int, seed: int, shift: int):
    """Rectified pair: right = left shifted by `shift` px (constant disparity) plus +-2 grey levels of independent noise."""
    left = synth_frame(width, height, seed)
    rng = np.random.RandomState(seed + 99)
    right = np.roll(left, -shift, axis=1).astype(np.int16) + rng.randint(-2, 3, left.shape)
    return left, np.clip(right, 0, 255).astype(np.uint8)


def synth_descriptors(n_query: int, n_db: int, seed: int = 7, planted_frac: float = 0.1):
    """Random 256-bit descriptors; planted_frac of the queries are database rows with 0..40 bit flips
    so the ratio test has accepts (config C4)."""
    rng = np.random.RandomState(seed)
    db = rng.randint(0, 256, (n_db, 32), dtype=np.uint8) if n_db else np.zeros((0, 32), np.uint8)
    q = rng.randint(0, 256, (n_query, 32), dtype=np.uint8)
    if n_db:
        npl = int(n_query * planted_frac)
        rows = rng.randint(0, n_db, npl)
        for i in range(npl):
            d = db[rows[i]].copy()
            nflip = rng.randint(0, 41)
            bits = rng.choice(256, nflip, replace=False)
            for b in bits:
                d[b >> 3] ^= np.uint8(1 << (b & 7))
            q[i] = d
    return q, db


def synth_guided_case(seed, n=600, m=500, w=752, h=480, radius=12.0, levels=8, dup=0.15, noise=6, skip=0.05, claims=0.9,
                      stereo=False, bounds=None, rot=20.0):
    """One (query set, target frame) scenario for the guided search (ORBmatcher::SearchByProjection): keypoints clustered
    like extractor output, queries = perturbed copies of keypoints (overlapping windows, several queries competing for
    one keypoint, a few bit errors per descriptor, 5 % unrelated descriptors). Returns a dict of arrays in the C-ABI
    layouts (kps KP_DTYPE [n], desc [n,32], queries QUERY_DTYPE [m], qdesc [m,32], bounds, uright or None, claimed [n])."""
    from ._lib import KP_DTYPE, QUERY_DTYPE
    rng = np.random.default_rng(seed)
    k = np.zeros(n, KP_DTYPE)
    centres = rng.uniform([0, 0], [w, h], (max(n // 12, 1), 2))
    c = centres[rng.integers(0, len(centres), n)]
    k["x"] = np.round(c[:, 0] + rng.normal(0, 14, n)).astype(np.float32)   # extractor output sits on (scaled) integer pixels
    k["y"] = np.round(c[:, 1] + rng.normal(0, 14, n)).astype(np.float32)
    pl = (1.0 / 1.2) ** np.arange(levels)   # keypoints per level fall off like mnFeaturesPerLevel (src/ORBextractor.cc:809-826)
    k["octave"] = rng.choice(levels, n, p=pl / pl.sum())
    k["angle"] = rng.uniform(0, 360, n).astype(np.float32)
    k["size"], k["response"], k["class_id"] = 31, 50, -1
    desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    src = rng.integers(0, max(n, 1), m)
    dups = rng.random(m) < dup   # several map points projecting onto the same keypoint
    src[dups] = src[rng.integers(0, max(m, 1), int(dups.sum()))]
    q = np.zeros(m, QUERY_DTYPE)
    q["u"] = k["x"][src] + rng.normal(0, 3, m).astype(np.float32)
    q["v"] = k["y"][src] + rng.normal(0, 3, m).astype(np.float32)
    q["radius"] = (radius * 1.2 ** k["octave"][src]).astype(np.float32)
    q["angle"] = (k["angle"][src] + rng.normal(rot, 4, m)).astype(np.float32) % np.float32(360)
    mode = rng.integers(0, 3, m)   # forward / backward / neutral level windows (src/ORBmatcher.cc:1774-1779)
    lo = np.where(mode == 0, k["octave"][src], np.where(mode == 1, 0, k["octave"][src] - 1))
    hi = np.where(mode == 0, -1, np.where(mode == 1, k["octave"][src], k["octave"][src] + 1))
    q["min_level"], q["max_level"] = lo, hi
    q["flags"] = np.where(rng.random(m) < skip, 1, 0) | np.where(rng.random(m) < claims, 2, 0)
    qdesc = desc[src].copy()
    flips = rng.integers(0, noise + 1, m)
    rows = np.arange(m)
    for j in range(noise):
        sel = flips > j
        b = rng.integers(0, 256, m)
        qdesc[rows[sel], (b >> 3)[sel]] ^= (1 << (b & 7)).astype(np.uint8)[sel]
    far = rng.random(m) < 0.05   # unrelated descriptors: bestDist > TH_HIGH
    qdesc[far] = rng.integers(0, 256, (int(far.sum()), 32), dtype=np.uint8)
    ur = None
    if stereo:
        ur = np.where(rng.random(n) < 0.7, k["x"] - rng.uniform(1, 40, n), -1).astype(np.float32)
        q["ur"] = np.where(rng.random(m) < 0.8, ur[src] + rng.normal(0, 4, m), q["u"] - 300).astype(np.float32)
    claimed = (rng.random(n) < 0.1).astype(np.uint8)
    return dict(kps=k, desc=desc, queries=q, qdesc=qdesc, bounds=bounds or (0.0, 0.0, float(w), float(h)), uright=ur, claimed=claimed)


def synth_vocabulary(k=10, L=3, seed=11, prune=0.05, stop=0.03):
    """Random DBoW2-style vocabulary tree in the node-table form of ORBvoc.txt (breadth-first ids): children = noisy copies of
    their parent (so descents are decided by few bits and ties occur), `prune` of the inner nodes end early as leaves or have
    fewer than k children, `stop` of the words have weight 0 (stopped words). Returns (k, L, parent, is_leaf, desc, weight)."""
    rng = np.random.default_rng(seed)
    parent, is_leaf, desc, weight = [0], [0], [np.zeros(32, np.uint8)], [0.0]
    frontier = [(0, 0, rng.integers(0, 256, 32, dtype=np.uint8))]   # (id, level, descriptor)
    while frontier:
        nxt = []
        for pid, lvl, pd in frontier:
            nch = k if rng.random() > prune else int(rng.integers(1, k + 1))
            for c in range(nch):
                d = pd.copy()
                flips = rng.integers(0, 256, int(rng.integers(0, 64 >> lvl) + 1))
                for b in flips:
                    d[b >> 3] ^= np.uint8(1 << (b & 7))
                if c > 0 and rng.random() < 0.03:
                    d = desc[-1].copy()   # identical siblings: equal distances, the first child must win
                nid = len(parent)
                leaf = lvl + 1 == L or (lvl + 1 >= 2 and rng.random() < prune)
                parent.append(pid); is_leaf.append(int(leaf)); desc.append(d)
                weight.append(0.0 if (leaf and rng.random() < stop) else float(rng.uniform(0.1, 9.0)) if leaf else 0.0)
                if not leaf:
                    nxt.append((nid, lvl + 1, d))
        frontier = nxt
    return k, L, np.array(parent, np.int32), np.array(is_leaf, np.uint8), np.stack(desc), np.array(weight, np.float64)


def synth_vocabulary_full(k=10, L=6, seed=5):
    """Balanced k^L-word vocabulary of the ORB vocabulary's shape (k = 10, L = 6: 1,111,110 nodes + root, 35 MB of descriptors),
    generated level by level without Python loops over nodes. Node ids breadth-first. Returns (k, L, parent, is_leaf, desc, weight)."""
    rng = np.random.default_rng(seed)
    parents, leaves, descs, weights = [np.zeros(1, np.int32)], [np.zeros(1, np.uint8)], [np.zeros((1, 32), np.uint8)], [np.zeros(1)]
    prev_ids = np.zeros(1, np.int64)
    prev_desc = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    next_id = 1
    for lvl in range(1, L + 1):
        n = len(prev_ids) * k
        par = np.repeat(prev_ids, k)
        d = np.repeat(prev_desc, k, axis=0)
        flips = rng.random((n, 256)) < (0.25 / lvl)   # children drift away from their parent, less at deeper levels
        d ^= np.packbits(flips, axis=1)
        ids = np.arange(next_id, next_id + n, dtype=np.int64)
        next_id += n
        parents.append(par.astype(np.int32)); descs.append(d)
        leaves.append(np.full(n, int(lvl == L), np.uint8))
        weights.append(rng.uniform(0.1, 9.0, n) if lvl == L else np.zeros(n))
        prev_ids, prev_desc = ids, d
    return k, L, np.concatenate(parents), np.concatenate(leaves), np.concatenate(descs), np.concatenate(weights)
