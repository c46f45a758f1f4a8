"""Development aid (GPU box): raw FAST keypoints of the CUDA path vs the oracle, per level, with the first differences printed.
Lives under tests/ because it uses the oracle. usage: python tests/fast_debug.py [case ...]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle_py as O
from orb_slam2_detailed_comments_b200 import ORBextractor, synth

CASES = {"tum": (640, 480, 1000, 1234), "kitti": (1241, 376, 2000, 1236), "euroc": (752, 480, 1200, 1237), "small": (200, 160, 500, 1238),
         "odd": (333, 257, 700, 1239), "4k": (3840, 2160, 8000, 1240)}


def run(name, ini=20, mn=7):
    w, h, nf, seed = CASES[name]
    img = synth.synth_frame(w, h, seed)
    ex = ORBextractor(nf, 1.2, 8, ini, mn)
    oe = O.OracleExtractor(nf, 1.2, 8, ini, mn)
    oe(img)
    try:
        ex.extract_batch(img[None])
    except Exception as e:  # capacity / TMA flags surface here
        print(name, "extract failed:", e)
    bad = 0
    for l in range(8):
        ro, rg = oe.raw_keypoints(l), ex.raw_keypoints(l)
        so = set(zip(ro["x"].astype(int), ro["y"].astype(int), ro["response"].astype(int)))
        sg = set(zip(rg["x"].astype(int), rg["y"].astype(int), rg["response"].astype(int)))
        same_order = len(ro) == len(rg) and all(np.array_equal(ro[f], rg[f]) for f in ("x", "y", "response"))
        pyr_ok = np.array_equal(ex.pyramid_level(l), oe.level(l))
        print(f"{name} L{l}: oracle {len(ro)} gpu {len(rg)} missing {len(so - sg)} extra {len(sg - so)} order_ok {same_order} pyramid_ok {pyr_ok}")
        if not same_order:
            bad += 1
            print("   missing (first 12):", sorted(so - sg)[:12])
            print("   extra   (first 12):", sorted(sg - so)[:12])
            if so == sg:
                i = next(i for i in range(len(ro)) if (ro["x"][i], ro["y"][i]) != (rg["x"][i], rg["y"][i]))
                print("   same set, first order difference at", i, ro[i - 1:i + 3][["x", "y", "response"]], rg[i - 1:i + 3][["x", "y", "response"]])
    return bad


if __name__ == "__main__":
    names = sys.argv[1:] or ["small", "odd", "tum", "kitti", "euroc"]
    tot = sum(run(n) for n in names)
    tot += run("tum", 40, 25) + run("small", 150, 130)
    print("levels with differences:", tot)
