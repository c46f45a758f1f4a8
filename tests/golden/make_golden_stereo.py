"""Generates tests/golden/stereo_*.npz: Frame::ComputeStereoMatches composed from REAL OpenCV calls (tests/cv2_composition.py
stereo_matches) on keypoints / descriptors / pyramids that the cv2 composition of the extractor produced. Pins the oracle's
orb_oracle_stereo_match (tests/test_oracle_golden.py) and through it the CUDA path.   python tests/golden/make_golden_stereo.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, ".."))

import cv2  # noqa: E402

import cv2_composition as comp  # noqa: E402
from orb_slam2_detailed_comments_b200 import synth  # noqa: E402

CASES = [  # name, width, height, nfeatures, seed, shift, bf, b
    ("stereo_480x360_nf600", 480, 360, 600, 311, 7, 60.0, 0.1),
    ("stereo_752x480_nf1200", 752, 480, 1200, 312, 12, 47.9, 0.11),
]


def main():
    cv2.setNumThreads(1)
    for name, w, h, nf, seed, shift, bf, b in CASES:
        left, right = synth.synth_stereo_pair(w, h, seed, shift)
        P = comp.Params(nf, 1.2, 8, 20, 7)
        kL, dL, lvL, _, _ = comp.extract(cv2, left, P, return_stages=True)
        kR, dR, lvR, _, _ = comp.extract(cv2, right, P, return_stages=True)
        ur, dp = comp.stereo_matches(cv2, kL, dL, lvL, kR, dR, lvR, P, bf, b)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), params=np.array([w, h, nf, seed, shift], np.int64),
                            bf=np.float32(bf), b=np.float32(b), uright=ur, depth=dp, n_left=len(kL), n_right=len(kR),
                            cv2_version=cv2.__version__)
        print(name, len(kL), "left keypoints,", int((ur >= 0).sum()), "matches")


if __name__ == "__main__":
    main()
