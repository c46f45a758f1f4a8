"""Generates tests/golden/*.npz with the REAL OpenCV (python cv2 4.13) composition of the reference pipeline
(tests/cv2_composition.py): every arithmetic primitive (resize, copyMakeBorder, FAST, GaussianBlur, fastAtan2) is
executed by OpenCV itself, only the reference's control flow is restated. Run in the build container:

    python tests/golden/make_golden.py

The fixtures pin the oracle (tests/test_oracle_golden.py) and, through it, the CUDA path. cv2 is NOT needed to run
the tests; only to regenerate these files."""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, ".."))

import cv2  # noqa: E402

import cv2_composition as comp  # noqa: E402
from orb_slam2_detailed_comments_b200 import synth  # noqa: E402

CASES = [
    # name, width, height, nfeatures, nlevels, scale, ini, min, seed
    ("tum_640x480_nf1000", 640, 480, 1000, 8, 1.2, 20, 7, 1234),
    ("kitti_1241x376_nf2000", 1241, 376, 2000, 8, 1.2, 20, 7, 1236),
    ("euroc_752x480_nf1200", 752, 480, 1200, 8, 1.2, 20, 7, 1237),
    ("small_320x240_nf500", 320, 240, 500, 8, 1.2, 20, 7, 42),
    ("small_400x300_nf300_l4_s15_th12", 400, 300, 300, 4, 1.5, 12, 7, 43),
]


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    cv2.setNumThreads(1)
    for name, w, h, nf, nl, sf, ini, mn, seed in CASES:
        img = synth.synth_frame(w, h, seed)
        P = comp.Params(nf, sf, nl, ini, mn)
        kps, desc, levels, blurred, raw = comp.extract(cv2, img, P, return_stages=True)
        out = {
            "params": np.array([w, h, nf, nl, ini, mn, seed], np.int64), "scale": np.float64(sf),
            "image_sha": sha(img),
            "kps": kps.astype(np.float32),          # x, y, size, angle, response, octave
            "desc": desc,
            "level_sha": np.array([sha(l) for l in levels]),
            "blur_sha": np.array([sha(b) if b is not None else "" for b in blurred]),
            "raw_counts": np.array([len(r) for r in raw], np.int64),
            "raw_sha": np.array([sha(np.array(r, np.float32).reshape(-1, 3)) for r in raw]),
            "cv2_version": cv2.__version__,
        }
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, len(kps), "keypoints")
    # primitive vectors: fastAtan2 on a grid, FAST on random patches
    rng = np.random.RandomState(5)
    y = rng.uniform(-3e6, 3e6, 4000).astype(np.float32)
    x = rng.uniform(-3e6, 3e6, 4000).astype(np.float32)
    y[:8] = [0, 0, 1, -1, 1, -1, 0, 5]
    x[:8] = [0, 1, 0, 0, 1, -1, -1, 5]
    ang = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    np.savez_compressed(os.path.join(HERE, "fast_atan2.npz"), y=y, x=x, angle=ang)
    src = synth.synth_frame(333, 257, 99)
    np.savez_compressed(os.path.join(HERE, "primitives_333x257.npz"), src=src,
                        resize=cv2.resize(src, (278, 214), interpolation=cv2.INTER_LINEAR),
                        border=cv2.copyMakeBorder(src, 19, 19, 19, 19, cv2.BORDER_REFLECT_101),
                        blur=cv2.GaussianBlur(src, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101),
                        fast20=np.array([(k.pt[0], k.pt[1], k.response) for k in cv2.FastFeatureDetector_create(20, True).detect(src)], np.float32),
                        fast7=np.array([(k.pt[0], k.pt[1], k.response) for k in cv2.FastFeatureDetector_create(7, True).detect(src)], np.float32))


if __name__ == "__main__":
    main()
