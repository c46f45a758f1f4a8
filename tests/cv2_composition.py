"""The reference pipeline composed from REAL OpenCV calls (python cv2), used to pin the oracle.

Second, independent restatement of /root/reference src/ORBextractor.cc in Python: every arithmetic
primitive the reference delegates to OpenCV is executed by the real library here (cv2.resize,
cv2.copyMakeBorder, cv2.FastFeatureDetector, cv2.GaussianBlur, cv2.fastAtan2); only control flow
(cell grid :1424-1601, quadtree :1050-1417, moments :104-161, rBRIEF sampling :177-254) is restated.
Slow (pure-Python loops) — only for golden generation and small pinned cases.
"""
from __future__ import annotations

import math
import os

import numpy as np

EDGE, PATCH, HALF = 19, 31, 15


def load_pattern():
    here = os.path.dirname(os.path.abspath(__file__))
    txt = open(os.path.join(here, "..", "include", "orb_pattern_31.inc")).read()
    vals = [int(v) for line in txt.splitlines() if not line.strip().startswith("//") for v in line.split(",") if v.strip()]
    return np.array(vals, np.int32).reshape(256, 4)


def cv_round(x) -> int:
    return int(np.rint(np.float32(x)))  # round-half-even on the float value


class Params:
    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        f32 = np.float32
        self.nfeatures, self.nlevels, self.ini_th, self.min_th = nfeatures, nlevels, ini_th, min_th
        sf = float(f32(scale_factor))  # ctor arg is float, member is double
        self.scale = [f32(1.0)]
        for _ in range(1, nlevels):
            self.scale.append(f32(float(self.scale[-1]) * sf))
        self.inv_scale = [f32(1.0) / s for s in self.scale]
        factor = f32(1.0 / sf)
        per = f32(f32(nfeatures) * (f32(1) - factor)) / (f32(1) - f32(math.pow(float(factor), float(nlevels))))
        self.quota, tot = [], 0
        for _ in range(nlevels - 1):
            q = cv_round(per)
            self.quota.append(q)
            tot += q
            per = f32(per * factor)
        self.quota.append(max(nfeatures - tot, 0))
        self.umax = [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]


def compute_pyramid(cv2, image, P: Params):
    levels = []
    for l in range(P.nlevels):
        s = P.inv_scale[l]
        w = cv_round(np.float32(image.shape[1]) * s)
        h = cv_round(np.float32(image.shape[0]) * s)
        if l == 0:
            temp = cv2.copyMakeBorder(image, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_REFLECT_101)
        else:
            inner = cv2.resize(levels[l - 1], (w, h), interpolation=cv2.INTER_LINEAR)  # from the PADDED level (F1)
            temp = cv2.copyMakeBorder(inner, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_REFLECT_101)
        levels.append(temp)
    return levels


def sincos_canonical(angle_f32):
    """Same operation sequence as the oracle/device (IEEE double, no FMA in CPython)."""
    x = float(angle_f32)
    invpio2, p1, p1t = 6.36619772367581382433e-01, 1.57079632673412561417e+00, 6.07710050650619224932e-11
    kf = math.floor(x * invpio2 + 0.5)
    r = (x - kf * p1) - kf * p1t
    z = r * r
    S1, S2, S3, S4, S5, S6 = (-1.66666666666666324348e-01, 8.33333333332248946124e-03, -1.98412698298579493134e-04,
                              2.75573137070700676789e-06, -2.50507602534068634195e-08, 1.58969099521155010221e-10)
    C1, C2, C3, C4, C5, C6 = (4.16666666666666019037e-02, -1.38888888888741095749e-03, 2.48015872894767294178e-05,
                              -2.75573143513906633035e-07, 2.08757232129817482790e-09, -1.13596475577881948265e-11)
    ps = S2 + z * (S3 + z * (S4 + z * (S5 + z * S6)))
    sn = r + (z * r) * (S1 + z * ps)
    pc = z * (C1 + z * (C2 + z * (C3 + z * (C4 + z * (C5 + z * C6)))))
    cs = 1.0 - (0.5 * z - z * pc)
    k = int(kf) & 3
    s, c = [(sn, cs), (cs, -sn), (-sn, -cs), (-cs, sn)][k]
    return np.float32(s), np.float32(c)


def distribute(keys, minX, maxX, minY, maxY, N):
    """keys: list of (x, y, response). Returns kept list in the reference's final list order.
    Canonical tie-break (size, creation seq), see DESIGN.md."""
    f32 = np.float32
    nIni = int(round(float(f32(maxX - minX) / f32(maxY - minY))))
    hX = f32(maxX - minX) / f32(nIni)
    nodes = []  # python list used as std::list: index 0 = front

    class Node:
        __slots__ = ("ulx", "uly", "brx", "bry", "keys", "nomore")

    ini = []
    for i in range(nIni):
        n = Node()
        n.ulx, n.uly = int(hX * f32(i)), 0
        n.brx, n.bry = int(hX * f32(i + 1)), maxY - minY
        n.keys, n.nomore = [], False
        nodes.append(n)
        ini.append(n)
    for k in keys:
        ini[int(f32(k[0]) / hX)].keys.append(k)
    nodes = [n for n in nodes if n.keys]
    for n in nodes:
        n.nomore = len(n.keys) == 1

    def divide(n):
        halfX = int(math.ceil(float(f32(n.brx - n.ulx) / f32(2))))
        halfY = int(math.ceil(float(f32(n.bry - n.uly) / f32(2))))
        mx, my = n.ulx + halfX, n.uly + halfY
        boxes = [(n.ulx, n.uly, mx, my), (mx, n.uly, n.brx, my), (n.ulx, my, mx, n.bry), (mx, my, n.brx, n.bry)]
        kids = []
        for b in boxes:
            c = Node()
            c.ulx, c.uly, c.brx, c.bry = b
            c.keys, c.nomore = [], False
            kids.append(c)
        for k in n.keys:
            kids[(0 if k[0] < mx else 1) + (0 if k[1] < my else 2)].keys.append(k)
        for c in kids:
            c.nomore = len(c.keys) == 1
        return kids

    finish = False
    seq = 0
    while not finish:
        prev_size = len(nodes)
        n_expand = 0
        cands = []
        new_front = []  # children pushed to the front, latest first
        survivors = []
        for n in nodes:
            if n.nomore:
                survivors.append(n)
                continue
            for c in divide(n):
                if c.keys:
                    new_front.insert(0, c)
                    if len(c.keys) > 1:
                        n_expand += 1
                        cands.append((len(c.keys), seq, c))
                        seq += 1
        nodes = new_front + survivors
        if len(nodes) >= N or len(nodes) == prev_size:
            finish = True
        elif len(nodes) + 3 * n_expand > N:
            while not finish:
                prev_size = len(nodes)
                prev = sorted(cands, key=lambda t: (t[0], t[1]))
                cands = []
                for size, _, node in reversed(prev):
                    for c in divide(node):
                        if c.keys:
                            nodes.insert(0, c)
                            if len(c.keys) > 1:
                                cands.append((len(c.keys), seq, c))
                                seq += 1
                    nodes.remove(node)
                    if len(nodes) >= N:
                        break
                if len(nodes) >= N or len(nodes) == prev_size:
                    finish = True
    out = []
    for n in nodes:
        best = n.keys[0]
        for k in n.keys[1:]:
            if k[2] > best[2]:
                best = k
        out.append(best)
    return out


def extract(cv2, image, P: Params, pattern=None, return_stages=False):
    f32 = np.float32
    pattern = load_pattern() if pattern is None else pattern
    levels = compute_pyramid(cv2, image, P)
    det_ini = cv2.FastFeatureDetector_create(P.ini_th, True)
    det_min = cv2.FastFeatureDetector_create(P.min_th, True)
    all_kps, raw_all = [], []
    for l, img in enumerate(levels):
        minB = EDGE - 3
        maxBX, maxBY = img.shape[1] - EDGE + 3, img.shape[0] - EDGE + 3
        width, height = f32(maxBX - minB), f32(maxBY - minB)
        nCols, nRows = int(width / f32(30)), int(height / f32(30))
        wCell, hCell = int(math.ceil(float(width / f32(nCols)))), int(math.ceil(float(height / f32(nRows))))
        raw = []
        for i in range(nRows):
            iniY = minB + i * hCell
            maxY = iniY + hCell + 6
            if iniY >= maxBY - 3:
                continue
            maxY = min(maxY, maxBY)
            for j in range(nCols):
                iniX = minB + j * wCell
                maxX = iniX + wCell + 6
                if iniX >= maxBX - 6:
                    continue
                maxX = min(maxX, maxBX)
                cell = img[iniY:maxY, iniX:maxX]
                k = det_ini.detect(cell)
                if not k:
                    k = det_min.detect(cell)
                for kp in k:
                    raw.append((kp.pt[0] + j * wCell, kp.pt[1] + i * hCell, kp.response))
        raw_all.append(raw)
        kept = distribute(raw, minB, maxBX, minB, maxBY, P.quota[l])
        size = float(int(f32(PATCH) * P.scale[l]))
        all_kps.append([[x + minB, y + minB, size, -1.0, r, l] for (x, y, r) in kept])
    # orientation on the un-blurred level
    for l, img in enumerate(levels):
        I = img.astype(np.int64)
        for kp in all_kps[l]:
            cx, cy = cv_round(kp[0]), cv_round(kp[1])
            m10 = m01 = 0
            for v in range(-HALF, HALF + 1):
                d = P.umax[abs(v)]
                row = I[cy + v, cx - d:cx + d + 1]
                u = np.arange(-d, d + 1)
                m10 += int((u * row).sum())
                m01 += v * int(row.sum())
            kp[3] = float(cv2.fastAtan2(float(f32(m01)), float(f32(m10))))
    # blur + descriptors + rescale
    out_k, out_d, blurred = [], [], []
    factorPI = f32(math.pi / 180.0)
    for l, img in enumerate(levels):
        if not all_kps[l]:
            blurred.append(None)
            continue
        b = cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        blurred.append(b)
        for kp in all_kps[l]:
            ang = f32(kp[3]) * factorPI
            sn, cs = sincos_canonical(ang)
            cx, cy = cv_round(kp[0]), cv_round(kp[1])
            px0, py0 = pattern[:, 0].astype(f32), pattern[:, 1].astype(f32)
            px1, py1 = pattern[:, 2].astype(f32), pattern[:, 3].astype(f32)

            def sample(px, py):
                yy = np.rint(px * sn + py * cs).astype(np.int64)  # f32 ops, separate roundings (no FMA)
                xx = np.rint(px * cs - py * sn).astype(np.int64)
                return b[cy + yy, cx + xx]

            bits = (sample(px0, py0) < sample(px1, py1)).astype(np.uint8)
            out_d.append(np.packbits(bits, bitorder="little"))
            k = list(kp)
            if l != 0:
                k[0] = float(f32(k[0]) * P.scale[l])
                k[1] = float(f32(k[1]) * P.scale[l])
            out_k.append(k)
    kps = np.array(out_k, np.float64).reshape(-1, 6)
    desc = np.array(out_d, np.uint8).reshape(-1, 32)
    if return_stages:
        return kps, desc, levels, blurred, raw_all
    return kps, desc


def stereo_matches(cv2, kL, dL, levelsL, kR, dR, levelsR, P: Params, bf, b):
    """Frame::ComputeStereoMatches (src/Frame.cc:880-1176) with the reference's control flow and float32 arithmetic; the
    patch arithmetic (convertTo CV_32F, centre subtraction, cv::norm NORM_L1) and the Hamming distance (cv2.norm NORM_HAMMING)
    are executed by the real OpenCV. kL / kR: rows (x, y, size, angle, response, octave). Returns (mvuRight, mvDepth) float32.
    Canonical choices as in the oracle: rows outside the level-0 image hold no candidates; an empty match list skips the filter."""
    f32 = np.float32
    N, Nr = len(kL), len(kR)
    uright = np.full(N, -1.0, f32)
    depth = np.full(N, -1.0, f32)
    TH_HIGH, TH_LOW = 100, 50
    thOrbDist = (TH_HIGH + TH_LOW) // 2
    nRows = levelsL[0].shape[0]
    rows = [[] for _ in range(nRows)]
    for iR in range(Nr):
        kpY = f32(kR[iR][1])
        r = f32(2.0) * P.scale[int(kR[iR][5])]
        maxr = int(math.ceil(float(f32(kpY + r))))
        minr = int(math.floor(float(f32(kpY - r))))
        for yi in range(minr, maxr + 1):
            if 0 <= yi < nRows:
                rows[yi].append(iR)
    bf, b = f32(bf), f32(b)
    minD, maxD = f32(0), f32(bf / b)
    dist_idx = []
    for iL in range(N):
        uL, vL, levelL = f32(kL[iL][0]), f32(kL[iL][1]), int(kL[iL][5])
        row = int(vL)
        if not (0 <= row < nRows) or not rows[row]:
            continue
        minU, maxU = f32(uL - maxD), f32(uL - minD)
        if maxU < 0:
            continue
        best, bestR = TH_HIGH, 0
        for iR in rows[row]:
            octR = int(kR[iR][5])
            if octR < levelL - 1 or octR > levelL + 1:
                continue
            uR = f32(kR[iR][0])
            if minU <= uR <= maxU:
                d = int(cv2.norm(dL[iL:iL + 1], dR[iR:iR + 1], cv2.NORM_HAMMING))
                if d < best:
                    best, bestR = d, iR
        if best >= thOrbDist:
            continue
        uR0 = f32(kR[bestR][0])
        sf = P.inv_scale[levelL]
        rnd = lambda v: f32(math.floor(float(v) + 0.5)) if v >= 0 else f32(-math.floor(-float(v) + 0.5))   # std::round
        suL, svL, suR0 = rnd(f32(uL * sf)), rnd(f32(vL * sf)), rnd(f32(uR0 * sf))
        w, L = 5, 5
        imL, imR = levelsL[levelL], levelsR[levelL]
        cy, cxl, cxr = int(svL), int(suL), int(suR0)
        IL = imL[cy - w:cy + w + 1, cxl - w:cxl + w + 1].astype(f32)
        IL = IL - IL[w, w] * np.ones(IL.shape, f32)
        iniu = f32(f32(suR0 - f32(L)) - f32(w))
        endu = f32(f32(f32(suR0 + f32(L)) + f32(w)) + f32(1))
        if iniu < 0 or endu >= imR.shape[1]:
            continue
        best_sad, best_inc, vd = 2 ** 31 - 1, 0, [f32(0)] * (2 * L + 1)
        for inc in range(-L, L + 1):
            IR = imR[cy - w:cy + w + 1, cxr + inc - w:cxr + inc + w + 1].astype(f32)
            IR = IR - IR[w, w] * np.ones(IR.shape, f32)
            dist = f32(cv2.norm(IL, IR, cv2.NORM_L1))
            if dist < f32(best_sad):
                best_sad, best_inc = int(dist), inc
            vd[L + inc] = dist
        if best_inc == -L or best_inc == L:
            continue
        d1, d2, d3 = vd[L + best_inc - 1], vd[L + best_inc], vd[L + best_inc + 1]
        with np.errstate(divide="ignore", invalid="ignore"):
            delta = f32(f32(d1 - d3) / f32(f32(2.0) * f32(f32(d1 + d3) - f32(f32(2.0) * d2))))
        if delta < -1 or delta > 1:
            continue
        best_uR = f32(P.scale[levelL] * f32(f32(suR0 + f32(best_inc)) + delta))
        disparity = f32(uL - best_uR)
        if disparity >= minD and disparity < maxD:
            if disparity <= 0:
                disparity = f32(0.01)
                best_uR = f32(float(uL) - 0.01)
            depth[iL] = f32(bf / disparity)
            uright[iL] = best_uR
            dist_idx.append((best_sad, iL))
    if dist_idx:
        dist_idx.sort()
        median = f32(dist_idx[len(dist_idx) // 2][0])
        th = f32(f32(f32(1.5) * f32(1.4)) * median)
        for d, i in reversed(dist_idx):
            if f32(d) < th:
                break
            uright[i] = -1
            depth[i] = -1
    return uright, depth
