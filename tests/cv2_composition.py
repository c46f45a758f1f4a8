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
