"""Level geometry and ALGORITHMIC byte counts of the extraction path (DESIGN.md "Roofline").

Pure arithmetic with the reference's types (float32 tables, src/ORBextractor.cc:800-811, 2099-2102); used by
bench.py for the roofline numerators and by the tests for shape checks. No compute happens here."""
from __future__ import annotations

import numpy as np

EDGE = 19


def scale_factors(nlevels: int = 8, scale_factor: float = 1.2) -> np.ndarray:
    sf = float(np.float32(scale_factor))          # ctor argument is float, member `scaleFactor` is double
    out = [np.float32(1.0)]
    for _ in range(1, nlevels):
        out.append(np.float32(float(out[-1]) * sf))
    return np.array(out, np.float32)


def level_sizes(width: int, height: int, nlevels: int = 8, scale_factor: float = 1.2):
    """[(interior w, interior h, padded cols, padded rows)] per level (padded = +38, this fork's mvImagePyramid)."""
    inv = np.float32(1.0) / scale_factors(nlevels, scale_factor)
    out = []
    for l in range(nlevels):
        w = int(np.rint(np.float32(width) * inv[l]))
        h = int(np.rint(np.float32(height) * inv[l]))
        out.append((w, h, w + 2 * EDGE, h + 2 * EDGE))
    return out


def stage_bytes(width: int, height: int, nfeatures: int, nlevels: int = 8, scale_factor: float = 1.2) -> dict:
    """Algorithmic bytes per image of each stage (what the stage must read + write once, no re-reads)."""
    P = [c * r for (_, _, c, r) in level_sizes(width, height, nlevels, scale_factor)]
    sumP = sum(P)
    return {
        "pyramid": width * height + sum(P[:-1]) + sumP,   # read input, read resize sources, write all levels
        "fast_cells": sumP,                               # read every level once
        "blur": 2 * sumP,                                 # read level, write blurred copy
        "quadtree": 0,                                    # control work on a few KB of packed keypoints
        "describe": 60 * nfeatures,                       # write 28-B keypoint + 32-B descriptor
    }


def algorithmic_bytes(width: int, height: int, nfeatures: int, nlevels: int = 8, scale_factor: float = 1.2) -> int:
    """A of SURVEY.md §8(d): C1 6,119,936; C2 9,224,499; C3 7,112,769; C5 141,758,480."""
    return sum(stage_bytes(width, height, nfeatures, nlevels, scale_factor).values())
