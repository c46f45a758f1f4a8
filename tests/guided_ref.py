"""Independent pure-Python restatement (float32 arithmetic via numpy scalars) of the guided-search pieces, used to
cross-check the C++ oracle: Frame::PosInGrid / AssignFeaturesToGrid / GetFeaturesInArea (src/Frame.cc:432-460, 633-745),
the SearchByProjection candidate loops (src/ORBmatcher.cc:69-184, 1702-1866) and ComputeThreeMaxima (:2026-2068).
Test infrastructure only."""
import math

import numpy as np

f32 = np.float32
COLS, ROWS, HISTO = 64, 48, 30


def _round_away(v) -> int:   # C round(): half away from zero (exact in double for a float32 input)
    v = float(v)
    return int(math.copysign(math.floor(abs(v) + 0.5), v))


class Grid:
    def __init__(self, kps, bounds):
        self.min_x, self.min_y = f32(bounds[0]), f32(bounds[1])
        self.inv_w = f32(COLS) / f32(f32(bounds[2]) - f32(bounds[0]))
        self.inv_h = f32(ROWS) / f32(f32(bounds[3]) - f32(bounds[1]))
        self.cells = {}
        for i in range(len(kps)):
            px = _round_away(f32(f32(kps["x"][i]) - self.min_x) * self.inv_w)
            py = _round_away(f32(f32(kps["y"][i]) - self.min_y) * self.inv_h)
            if 0 <= px < COLS and 0 <= py < ROWS:
                self.cells.setdefault((px, py), []).append(i)
        self.kps = kps

    def in_area(self, x, y, r, min_level=-1, max_level=-1):
        x, y, r = f32(x), f32(y), f32(r)
        cx0 = max(0, math.floor(float(f32(f32(x - self.min_x) - r) * self.inv_w)))
        if cx0 >= COLS:
            return []
        cx1 = min(COLS - 1, math.ceil(float(f32(f32(x - self.min_x) + r) * self.inv_w)))
        if cx1 < 0:
            return []
        cy0 = max(0, math.floor(float(f32(f32(y - self.min_y) - r) * self.inv_h)))
        if cy0 >= ROWS:
            return []
        cy1 = min(ROWS - 1, math.ceil(float(f32(f32(y - self.min_y) + r) * self.inv_h)))
        if cy1 < 0:
            return []
        check = min_level > 0 or max_level >= 0
        out = []
        k = self.kps
        for ix in range(cx0, cx1 + 1):
            for iy in range(cy0, cy1 + 1):
                for i in self.cells.get((ix, iy), ()):
                    if check:
                        if k["octave"][i] < min_level:
                            continue
                        if max_level >= 0 and k["octave"][i] > max_level:
                            continue
                    if abs(f32(k["x"][i]) - x) < r and abs(f32(k["y"][i]) - y) < r:
                        out.append(i)
        return out


def three_maxima(sizes):
    ind1 = ind2 = ind3 = -1
    max1 = max2 = max3 = 0
    for i, s in enumerate(sizes):
        if s > max1:
            max3, max2, max1 = max2, max1, s
            ind3, ind2, ind1 = ind2, ind1, i
        elif s > max2:
            max3, max2 = max2, s
            ind3, ind2 = ind2, i
        elif s > max3:
            max3, ind3 = s, i
    if f32(max2) < f32(0.1) * f32(max1):
        ind2 = ind3 = -1
    elif f32(max3) < f32(0.1) * f32(max1):
        ind3 = -1
    return ind1, ind2, ind3


def search_by_projection(kps, desc, queries, qdesc, bounds, th_dist=100, nn_ratio=0.0, check_orientation=True, uright=None,
                         claimed=None):
    grid = Grid(kps, bounds)
    n, m = len(kps), len(queries)
    cl = np.zeros(n, bool) if claimed is None else np.asarray(claimed).astype(bool).copy()
    assigned = np.full(n, -1, np.int32)
    qmatch = np.full(m, -1, np.int32)
    hist = [[] for _ in range(HISTO)]
    factor = f32(HISTO) / f32(360.0)
    nm = 0
    bits = np.unpackbits(desc, axis=1) if n else np.zeros((0, 256), np.uint8)
    qbits = np.unpackbits(qdesc, axis=1) if m else np.zeros((0, 256), np.uint8)
    for iq in range(m):
        Q = queries[iq]
        if Q["flags"] & 1:
            continue
        cand = grid.in_area(Q["u"], Q["v"], Q["radius"], int(Q["min_level"]), int(Q["max_level"]))
        best = best2 = 256
        lvl = lvl2 = -1
        bidx = -1
        for i in cand:
            if cl[i]:
                continue
            if uright is not None and uright[i] > 0:
                if abs(f32(Q["ur"]) - f32(uright[i])) > f32(Q["radius"]):
                    continue
            d = int(np.count_nonzero(bits[i] != qbits[iq]))
            if d < best:
                best2, best = best, d
                lvl2, lvl = lvl, int(kps["octave"][i])
                bidx = i
            elif d < best2:
                lvl2, best2 = int(kps["octave"][i]), d
        if bidx < 0 or best > th_dist:
            continue
        if nn_ratio > 0 and lvl == lvl2 and f32(best) > f32(nn_ratio) * f32(best2):
            continue
        assigned[bidx] = iq
        qmatch[iq] = bidx
        if Q["flags"] & 2:
            cl[bidx] = True
        nm += 1
        if check_orientation:
            rot = f32(Q["angle"]) - f32(kps["angle"][bidx])
            if rot < 0:
                rot = f32(rot + f32(360.0))
            b = _round_away(f32(rot * factor))
            if b == HISTO:
                b = 0
            if 0 <= b < HISTO:
                hist[b].append((iq, bidx))
    if check_orientation:
        keep = three_maxima([len(h) for h in hist])
        for b in range(HISTO):
            if b in keep:
                continue
            for iq, t in hist[b]:
                assigned[t] = -1
                qmatch[iq] = -1
                nm -= 1
    return assigned, qmatch, nm


def search_for_initialization(kps, desc, queries, qdesc, bounds, th_dist=50, nn_ratio=0.9, check_orientation=True):
    """src/ORBmatcher.cc:570-712, independent Python statement."""
    grid = Grid(kps, bounds)
    n, m = len(kps), len(queries)
    m12 = np.full(m, -1, np.int32)
    m21 = np.full(n, -1, np.int32)
    mdist = np.full(n, 2 ** 31 - 1, np.int64)
    hist = [[] for _ in range(HISTO)]
    factor = f32(HISTO) / f32(360.0)
    bits = np.unpackbits(desc, axis=1) if n else np.zeros((0, 256), np.uint8)
    qbits = np.unpackbits(qdesc, axis=1) if m else np.zeros((0, 256), np.uint8)
    nm = 0
    for i1 in range(m):
        Q = queries[i1]
        if Q["flags"] & 1:
            continue
        cand = grid.in_area(Q["u"], Q["v"], Q["radius"], int(Q["min_level"]), int(Q["max_level"]))
        best = best2 = 2 ** 31 - 1
        bidx = -1
        for i2 in cand:
            d = int(np.count_nonzero(bits[i2] != qbits[i1]))
            if mdist[i2] <= d:
                continue
            if d < best:
                best2, best, bidx = best, d, i2
            elif d < best2:
                best2 = d
        if best <= th_dist and f32(best) < f32(f32(best2) * f32(nn_ratio)):
            if m21[bidx] >= 0:
                m12[m21[bidx]] = -1
                nm -= 1
            m12[i1], m21[bidx], mdist[bidx] = bidx, i1, best
            nm += 1
            if check_orientation:
                rot = f32(Q["angle"]) - f32(kps["angle"][bidx])
                if rot < 0:
                    rot = f32(rot + f32(360.0))
                b = _round_away(f32(rot * factor))
                if b == HISTO:
                    b = 0
                if 0 <= b < HISTO:
                    hist[b].append(i1)
    if check_orientation:
        keep = three_maxima([len(h) for h in hist])
        for b in range(HISTO):
            if b in keep:
                continue
            for i1 in hist[b]:
                if m12[i1] >= 0:
                    m12[i1] = -1
                    nm -= 1
    inv = np.full(n, -1, np.int32)
    for i1 in range(m):
        if m12[i1] >= 0:
            inv[m12[i1]] = i1
    return m12, inv, nm
