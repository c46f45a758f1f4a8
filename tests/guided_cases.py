"""Seeded (query set, target frame) scenarios for the guided-search tests: keypoints clustered like extractor output,
queries = perturbed copies of the keypoints so that windows overlap and several queries compete for one keypoint."""
import numpy as np

from oracle.oracle_py import QUERY_DTYPE

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])


def make_case(seed, n=600, m=500, w=752, h=480, radius=12.0, levels=8, dup=0.15, noise=6, skip=0.05, claims=0.9, stereo=False,
              bounds=None, rot=20.0):
    rng = np.random.default_rng(seed)
    k = np.zeros(n, KP_DTYPE)
    centres = rng.uniform([0, 0], [w, h], (max(n // 12, 1), 2))
    c = centres[rng.integers(0, len(centres), n)]
    k["x"] = np.round(c[:, 0] + rng.normal(0, 14, n)).astype(np.float32)   # extractor output sits on (scaled) integer pixels
    k["y"] = np.round(c[:, 1] + rng.normal(0, 14, n)).astype(np.float32)
    k["octave"] = rng.integers(0, levels, n)
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
    mode = rng.integers(0, 3, m)   # forward / backward / neutral level windows (src/ORBmatcher.cc:1774-1779) and the local-map window
    lo = np.where(mode == 0, k["octave"][src], np.where(mode == 1, 0, k["octave"][src] - 1))
    hi = np.where(mode == 0, -1, np.where(mode == 1, k["octave"][src], k["octave"][src] + 1))
    q["min_level"], q["max_level"] = lo, hi
    q["flags"] = np.where(rng.random(m) < skip, 1, 0) | np.where(rng.random(m) < claims, 2, 0)
    qdesc = desc[src].copy()
    flips = rng.integers(0, noise + 1, m)   # a few bit errors per query; some queries far away (bestDist > TH_HIGH)
    far = rng.random(m) < 0.05
    flips[far] = 140
    for i in range(m):
        for b in rng.integers(0, 256, flips[i]):
            qdesc[i, b >> 3] ^= np.uint8(1 << (b & 7))
    ur = None
    if stereo:
        ur = np.where(rng.random(n) < 0.7, k["x"] - rng.uniform(1, 40, n), -1).astype(np.float32)
        q["ur"] = np.where(rng.random(m) < 0.8, ur[src] + rng.normal(0, 4, m), q["u"] - 300).astype(np.float32)
    claimed = (rng.random(n) < 0.1).astype(np.uint8)
    return dict(kps=k, desc=desc, queries=q, qdesc=qdesc, bounds=bounds or (0.0, 0.0, float(w), float(h)), uright=ur, claimed=claimed)
