"""Offline FAST statistics on synthetic frames (CPU, numpy + the oracle's score map): how many detection pixels pass each
quick-reject level, how many are corners, how many survive NMS. Drives the choice of reject tests in csrc/fast.cu
(profiles/r2_fast_phases.md). Analysis tool under tests/ because it uses the oracle (test infrastructure); never part of the product path."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle_py as O
from orb_slam2_detailed_comments_b200 import synth

RING = [(0, 3), (1, 3), (2, 2), (3, 1), (3, 0), (3, -1), (2, -2), (1, -3), (0, -3), (-1, -3), (-2, -2), (-3, -1), (-3, 0), (-3, 1), (-2, 2), (-1, 3)]


def shifted(img, dx, dy):
    h, w = img.shape
    return img[3 + dy:h - 3 + dy, 3 + dx:w - 3 + dx].astype(np.int16)


def stats(name, seed=1234):
    W, H, nf = synth.CONFIGS[name]
    ex = O.OracleExtractor(nf, 1.2, 8, 20, 7)
    img = synth.synth_frame(W, H, seed)
    ex(img)
    tot = dict(px=0)
    rows = []
    for l in range(8):
        lv = ex.level(l)
        # detection domain of the level: [19, dim-19) roughly (minBorder 16 + 3); use that interior
        sub = lv[16:lv.shape[0] - 16, 16:lv.shape[1] - 16]
        S = O.fast_score_map(sub)[3:-3, 3:-3].astype(np.int16)
        c = shifted(sub, 0, 0)
        d = [shifted(sub, dx, dy) - c for dx, dy in RING]   # ring - centre
        r = dict(level=l, px=c.size)
        for th, tag in ((20, "ini"), (7, "min")):
            a = [np.abs(x) > th for x in d]
            br = [x > th for x in d]
            dk = [x < -th for x in d]
            compass_abs = (a[0] | a[8]) & (a[4] | a[12])
            compass_sgn = ((br[0] | br[8]) & (br[4] | br[12])) | ((dk[0] | dk[8]) & (dk[4] | dk[12]))
            # 4 consecutive of the 8 even ring positions, same polarity
            def four(m):
                ev = [m[2 * k] for k in range(8)]
                out = np.zeros_like(ev[0])
                for k in range(8):
                    out |= ev[k] & ev[(k + 1) % 8] & ev[(k + 2) % 8] & ev[(k + 3) % 8]
                return out
            even4 = four(br) | four(dk)
            # 2 adjacent compass points of same polarity (any 9-arc holds two adjacent compass points)
            def adj2(m):
                cp = [m[0], m[4], m[8], m[12]]
                out = np.zeros_like(cp[0])
                for k in range(4):
                    out |= cp[k] & cp[(k + 1) % 4]
                return out
            comp_adj = adj2(br) | adj2(dk)
            r[tag + "_abs"] = int(compass_abs.sum())
            r[tag + "_sgn"] = int(compass_sgn.sum())
            r[tag + "_adj"] = int(comp_adj.sum())
            r[tag + "_ev4"] = int((even4).sum())
            r[tag + "_corner"] = int((S >= th).sum())
        rows.append(r)
    keys = [k for k in rows[0] if k != "level"]
    tot = {k: sum(r[k] for r in rows) for k in keys}
    print(name, "seed", seed)
    print("level " + " ".join(f"{k:>10}" for k in keys))
    for r in rows:
        print(f"{r['level']:5d} " + " ".join(f"{r[k]:10d}" for k in keys))
    print("total " + " ".join(f"{tot[k]:10d}" for k in keys))
    print("frac  " + " ".join(f"{tot[k] / tot['px']:10.4f}" for k in keys))


if __name__ == "__main__":
    for n in sys.argv[1:] or ["kitti"]:
        stats(n)
