#!/usr/bin/env python
"""Per-call latency of the extractor the way the reference's Tracking thread uses it (src/Frame.cc:468-481: one image per
ORBextractor::operator() call, or the L and R image of a stereo frame): wall clock around orb_extract_batch /
orb_extract_stereo_batch with host buffers in and out, batch 1 / 2 / 2 (stereo incl. ComputeStereoMatches), pageable and pinned.
Prints one JSON line. usage: python tools/latency_bench.py [--iters 300] [--config kitti|tum|euroc]"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CONFIGS = {"kitti": (1241, 376, 2000), "tum": (640, 480, 1000), "euroc": (752, 480, 1200)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=300)
    ap.add_argument("--config", default="kitti")
    args = ap.parse_args()
    import torch
    import orb_slam2_detailed_comments_b200 as pkg
    from orb_slam2_detailed_comments_b200 import synth
    from orb_slam2_detailed_comments_b200._lib import lib, check
    W, H, nF = CONFIGS[args.config]
    frames = synth.synth_batch(W, H, 16, seed0=1234, distinct=16)
    L = lib()
    out = {"config": f"{args.config} {W}x{H}, nFeatures={nF}, 8 levels, scale 1.2, FAST 20/7; wall clock per call, host buffers in and out",
           "iters": args.iters}
    for name, B, stereo in (("mono_1_image", 1, False), ("two_images", 2, False), ("stereo_pair_with_matches", 2, True)):
        ex = pkg.ORBextractor(nF, 1.2, 8, 20, 7)
        cap = ex.max_keypoints()
        for pinned in (True, False):
            mk = (lambda *s: torch.empty(s, dtype=torch.uint8).pin_memory().numpy()) if pinned else (lambda *s: np.empty(s, np.uint8))
            src = mk(16, H, W)
            src[:] = frames
            kps, desc = mk(B, cap, 28), mk(B, cap, 32)
            ur, dp = mk(cap * 4).view(np.float32), mk(cap * 4).view(np.float32)
            counts, matched = np.zeros(B, np.int32), np.zeros(1, np.int32)
            vp = lambda a: a.ctypes.data_as(C.c_void_p)

            def call(i):
                f0 = (i * B) % 16
                ptrs = (src.ctypes.data + (f0 + np.arange(B, dtype=np.uint64)) * np.uint64(W * H)).astype(np.uint64)
                if stereo:
                    check(L.orb_extract_stereo_batch(ex._h, vp(ptrs), 0, W, W, H, 1, vp(kps), vp(desc), 0, cap, vp(counts), None,
                                                     C.c_float(386.1448), C.c_float(0.537), vp(ur), vp(dp), vp(matched)))
                else:
                    check(L.orb_extract_batch(ex._h, vp(ptrs), 0, W, W, H, B, vp(kps), vp(desc), 0, cap, vp(counts), None))
            for i in range(20):
                call(i)
            ts = []
            for i in range(args.iters):
                t0 = time.perf_counter()
                call(i)
                ts.append(1e3 * (time.perf_counter() - t0))
            ts = np.sort(np.array(ts))
            out.setdefault(name, {})["pinned" if pinned else "pageable"] = {
                "median_ms": float(np.median(ts)), "p10_ms": float(ts[len(ts) // 10]), "p99_ms": float(ts[int(len(ts) * 0.99)]),
                "keypoints": int(counts.sum())}
        ex.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
